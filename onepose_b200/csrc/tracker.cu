// SURVEY.md 8(f) row f4 -- tracker-side geometry (src/tracker/ba_tracker.py, src/tracker/tracking_utils.py), float64 throughout.
//
//  k_project_points   : `project` (tracking_utils.py:74-89) + the mean reprojection distance of flow_track (ba_tracker.py:318-319)
//  k_triangulate      : `apply_triangulation` (ba_tracker.py:267-273) = cv2.triangulatePoints: per correspondence the 4x4 DLT matrix of
//                       :255-259, right singular vector of the smallest singular value (cyclic Jacobi on A^T A), de-homogenised
//  k_snavely_residual : `SnavelyReprojectionErrorV2` + `AngleAxisRotatePoint` (tracking_utils.py:91-169), with the gather by ptIdx / camIdx
//  op_ba_lm           : the windowed bundle adjustment of `apply_ba` (ba_tracker.py:358-441).  The reference hands the residual function to
//                       DeepLM's `Solve` (third-party, not vendored, unpinned: github.com/hjwdzh/DeepLM); here Levenberg-Marquardt on the
//                       same residual with analytic Jacobians, the point blocks eliminated by the Schur complement and the reduced camera
//                       system (6 x cameras) solved by a dense Cholesky in one CTA.  The whole iteration -- including accept / reject and
//                       the damping update -- runs on the device: nothing is copied to the host between iterations.
#include "common.cuh"

namespace opb {
namespace trk {

// ---------------------------------------------------------------------------------------------- small helpers
struct Obs { double r[2], Jc[2][6], Jp[2][3]; };

// R(w) p as AngleAxisRotatePoint (tracking_utils.py:91-140): Rodrigues for theta^2 > 0, p + w x p otherwise
__device__ __forceinline__ void rotate(const double *w, const double *p, double *q) {
    const double t2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    if (t2 > 0.0) {
        const double t = sqrt(t2), c = cos(t), s = sin(t), ti = 1.0 / t;
        const double k0 = w[0] * ti, k1 = w[1] * ti, k2 = w[2] * ti;
        const double x0 = k1 * p[2] - k2 * p[1], x1 = k2 * p[0] - k0 * p[2], x2 = k0 * p[1] - k1 * p[0];
        const double tmp = (k0 * p[0] + k1 * p[1] + k2 * p[2]) * (1.0 - c);
        q[0] = p[0] * c + x0 * s + k0 * tmp; q[1] = p[1] * c + x1 * s + k1 * tmp; q[2] = p[2] * c + x2 * s + k2 * tmp;
    } else {
        q[0] = p[0] + (w[1] * p[2] - w[2] * p[1]); q[1] = p[1] + (w[2] * p[0] - w[0] * p[2]); q[2] = p[2] + (w[0] * p[1] - w[1] * p[0]);
    }
}

// residual of one observation: cam = (angle-axis, translation), feat = (x, y, focal, l1, l2)  (tracking_utils.py:143-169)
__device__ __forceinline__ void residual(const double *cam, const double *pt, const double *feat, double *r) {
    double q[3];
    rotate(cam, pt, q);
    q[0] += cam[3]; q[1] += cam[4]; q[2] += cam[5];
    r[0] = feat[2] * (q[0] / q[2]) + feat[3] - feat[0];
    r[1] = feat[2] * (q[1] / q[2]) + feat[4] - feat[1];
}

// residual + analytic Jacobians.  d(Rp)/dw = -R [p]x Jr(w), Jr = I - (1-cos t)/t^2 [w]x + (t - sin t)/t^3 [w]x^2 (right Jacobian of SO(3));
// d(Rp)/dp = R.
__device__ void linearize(const double *cam, const double *pt, const double *feat, Obs &o) {
    const double *w = cam;
    const double t2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    double R[3][3], Jr[3][3];
    const double W[3][3] = {{0, -w[2], w[1]}, {w[2], 0, -w[0]}, {-w[1], w[0], 0}};
    double W2[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) W2[i][j] = W[i][0] * W[0][j] + W[i][1] * W[1][j] + W[i][2] * W[2][j];
    double a, b, c1, c2;                      // R = I + a W + b W^2,  Jr = I - c1 W + c2 W^2
    if (t2 > 1e-12) {
        const double t = sqrt(t2);
        a = sin(t) / t; b = (1.0 - cos(t)) / t2; c1 = b; c2 = (t - sin(t)) / (t2 * t);
    } else {
        a = 1.0 - t2 / 6.0; b = 0.5 - t2 / 24.0; c1 = b; c2 = 1.0 / 6.0 - t2 / 120.0;
        if (t2 == 0.0) b = 0.0;               // the reference's zero-angle branch is first order: p + w x p
    }
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            const double id = i == j ? 1.0 : 0.0;
            R[i][j] = id + a * W[i][j] + b * W2[i][j];
            Jr[i][j] = id - c1 * W[i][j] + c2 * W2[i][j];
        }
    double q[3];
    for (int i = 0; i < 3; ++i) q[i] = R[i][0] * pt[0] + R[i][1] * pt[1] + R[i][2] * pt[2] + cam[3 + i];
    const double f = feat[2], iz = 1.0 / q[2];
    o.r[0] = f * q[0] * iz + feat[3] - feat[0];
    o.r[1] = f * q[1] * iz + feat[4] - feat[1];
    const double dq[2][3] = {{f * iz, 0.0, -f * q[0] * iz * iz}, {0.0, f * iz, -f * q[1] * iz * iz}};
    // M = -R [p]x Jr
    const double P[3][3] = {{0, -pt[2], pt[1]}, {pt[2], 0, -pt[0]}, {-pt[1], pt[0], 0}};
    double RP[3][3], M[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) RP[i][j] = R[i][0] * P[0][j] + R[i][1] * P[1][j] + R[i][2] * P[2][j];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) M[i][j] = -(RP[i][0] * Jr[0][j] + RP[i][1] * Jr[1][j] + RP[i][2] * Jr[2][j]);
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 3; ++j) {
            o.Jc[i][j] = dq[i][0] * M[0][j] + dq[i][1] * M[1][j] + dq[i][2] * M[2][j];
            o.Jc[i][3 + j] = dq[i][j];
            o.Jp[i][j] = dq[i][0] * R[0][j] + dq[i][1] * R[1][j] + dq[i][2] * R[2][j];
        }
}

// ---------------------------------------------------------------------------------------------- stand-alone geometry kernels
__global__ void k_snavely_residual(const double *__restrict__ points, const double *__restrict__ cams, const double *__restrict__ feats,
                                   const int64_t *__restrict__ pt_idx, const int64_t *__restrict__ cam_idx, long long n, double *__restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *p = points + (pt_idx ? pt_idx[i] : i) * 3, *c = cams + (cam_idx ? cam_idx[i] : i) * 6;
    residual(c, p, feats + i * 5, out + i * 2);
}

__global__ void k_project_points(const double *__restrict__ xyz, const double *__restrict__ K, const double *__restrict__ RT, long long n,
                                 const double *__restrict__ ref2d, double *__restrict__ xy, double *__restrict__ depth, double *__restrict__ dist_sum) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double d = 0.0;
    if (i < n) {
        const double *p = xyz + i * 3;
        double q[3], h[3];
        for (int r = 0; r < 3; ++r) q[r] = p[0] * RT[r * 4] + p[1] * RT[r * 4 + 1] + p[2] * RT[r * 4 + 2] + RT[r * 4 + 3];
        for (int r = 0; r < 3; ++r) h[r] = q[0] * K[r * 3] + q[1] * K[r * 3 + 1] + q[2] * K[r * 3 + 2];
        const double x = h[0] / h[2], y = h[1] / h[2];
        xy[i * 2] = x; xy[i * 2 + 1] = y;
        if (depth) depth[i] = q[2];
        if (ref2d) { const double dx = x - ref2d[i * 2], dy = y - ref2d[i * 2 + 1]; d = sqrt(dx * dx + dy * dy); }
    }
    if (dist_sum) {
        d = warp_sum(d);
        if ((threadIdx.x & 31) == 0 && d != 0.0) atomicAdd(dist_sum, d);
    }
}

// smallest-eigenvalue eigenvector of the symmetric 4x4 matrix S (cyclic Jacobi, float64)
__device__ void smallest_eigvec4(double S[4][4], double *v) {
    double V[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0;
        for (int i = 0; i < 4; ++i)
            for (int j = i + 1; j < 4; ++j) off += S[i][j] * S[i][j];
        if (off < 1e-300) break;
        for (int p = 0; p < 3; ++p)
            for (int q = p + 1; q < 4; ++q) {
                if (S[p][q] == 0.0) continue;
                const double theta = (S[q][q] - S[p][p]) / (2.0 * S[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 4; ++k) { const double a = S[k][p], b = S[k][q]; S[k][p] = c * a - s * b; S[k][q] = s * a + c * b; }
                for (int k = 0; k < 4; ++k) { const double a = S[p][k], b = S[q][k]; S[p][k] = c * a - s * b; S[q][k] = s * a + c * b; }
                for (int k = 0; k < 4; ++k) { const double a = V[k][p], b = V[k][q]; V[k][p] = c * a - s * b; V[k][q] = s * a + c * b; }
            }
    }
    int m = 0;
    for (int i = 1; i < 4; ++i)
        if (S[i][i] < S[m][m]) m = i;
    for (int k = 0; k < 4; ++k) v[k] = V[k][m];
}

__global__ void k_triangulate(const double *__restrict__ P1, const double *__restrict__ P2, const double *__restrict__ x1,
                              const double *__restrict__ x2, long long n, double *__restrict__ X) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double A[4][4];
    for (int k = 0; k < 4; ++k) {                                           // ba_tracker.py:255-259
        A[0][k] = x1[i * 2] * P1[8 + k] - P1[k];
        A[1][k] = x1[i * 2 + 1] * P1[8 + k] - P1[4 + k];
        A[2][k] = x2[i * 2] * P2[8 + k] - P2[k];
        A[3][k] = x2[i * 2 + 1] * P2[8 + k] - P2[4 + k];
    }
    double S[4][4], v[4];
    for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b) S[a][b] = A[0][a] * A[0][b] + A[1][a] * A[1][b] + A[2][a] * A[2][b] + A[3][a] * A[3][b];
    smallest_eigvec4(S, v);
    X[i * 3] = v[0] / v[3]; X[i * 3 + 1] = v[1] / v[3]; X[i * 3 + 2] = v[2] / v[3];
}

// ---------------------------------------------------------------------------------------------- bundle adjustment (LM + Schur)
// state: [0] lambda, [1] cost at the current estimate, [2] cost at the candidate, [3] accepted steps, [4] done flag, [5] last step accepted,
//        [6] iterations run, [7] initial cost
struct Ba {
    int C, P, O, n;                   // cameras, points, observations, n = 6 C
    double *cams, *pts;               // current estimate [C,6], [P,3]
    double *cams_new, *pts_new;       // candidate
    const double *feats;              // [O,5]
    const int *cam_idx, *pt_idx;      // [O]
    const int *pt_off, *pt_obs;       // observations grouped by point: obs ids pt_obs[pt_off[p] .. pt_off[p+1])
    double *U, *V, *gc, *gp, *Wo, *Vinv, *S, *rhs, *dc, *state;
    int max_success;
};

__global__ void k_ba_zero(Ba b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (b.state[4] != 0.0) return;
    if (i < b.C * 36) b.U[i] = 0.0;
    if (i < b.P * 9) b.V[i] = 0.0;
    if (i < b.C * 6) b.gc[i] = 0.0;
    if (i < b.P * 3) b.gp[i] = 0.0;
    if (i < b.n * b.n) b.S[i] = 0.0;
    if (i == 0) { b.state[1] = 0.0; b.state[2] = 0.0; }
}

__global__ void k_ba_linearize(Ba b) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (b.state[4] != 0.0) return;
    double cost = 0.0;
    if (o < b.O) {
        const int c = b.cam_idx[o], p = b.pt_idx[o];
        Obs ob;
        linearize(b.cams + c * 6, b.pts + p * 3, b.feats + o * 5, ob);
        cost = ob.r[0] * ob.r[0] + ob.r[1] * ob.r[1];
        for (int i = 0; i < 6; ++i) {
            for (int j = 0; j < 6; ++j) atomicAdd(&b.U[c * 36 + i * 6 + j], ob.Jc[0][i] * ob.Jc[0][j] + ob.Jc[1][i] * ob.Jc[1][j]);
            atomicAdd(&b.gc[c * 6 + i], ob.Jc[0][i] * ob.r[0] + ob.Jc[1][i] * ob.r[1]);
            for (int j = 0; j < 3; ++j) b.Wo[(long long)o * 18 + i * 3 + j] = ob.Jc[0][i] * ob.Jp[0][j] + ob.Jc[1][i] * ob.Jp[1][j];
        }
        for (int i = 0; i < 3; ++i) {
            for (int j = 0; j < 3; ++j) atomicAdd(&b.V[p * 9 + i * 3 + j], ob.Jp[0][i] * ob.Jp[0][j] + ob.Jp[1][i] * ob.Jp[1][j]);
            atomicAdd(&b.gp[p * 3 + i], ob.Jp[0][i] * ob.r[0] + ob.Jp[1][i] * ob.r[1]);
        }
    }
    cost = warp_sum(cost);
    if ((threadIdx.x & 31) == 0 && cost != 0.0) atomicAdd(&b.state[1], cost);
}

// damped camera blocks onto the diagonal of S, rhs = -gc; damped point blocks inverted.  Damping: M + lambda * diag(M) (Marquardt).
__global__ void k_ba_blocks(Ba b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (b.state[4] != 0.0) return;
    const double lam = b.state[0];
    if (i < b.C) {
        for (int r = 0; r < 6; ++r) {
            for (int c = 0; c < 6; ++c) {
                double v = b.U[i * 36 + r * 6 + c];
                if (r == c) v += lam * v + 1e-12;
                atomicAdd(&b.S[(long long)(i * 6 + r) * b.n + i * 6 + c], v);
            }
            b.rhs[i * 6 + r] = -b.gc[i * 6 + r];
        }
    }
    if (i < b.P) {
        double m[3][3];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) m[r][c] = b.V[i * 9 + r * 3 + c];
        for (int r = 0; r < 3; ++r) m[r][r] += lam * m[r][r] + 1e-12;
        const double det = m[0][0] * (m[1][1] * m[2][2] - m[1][2] * m[2][1]) - m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) +
                           m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
        const double id = 1.0 / det;
        double *inv = b.Vinv + i * 9;
        inv[0] = (m[1][1] * m[2][2] - m[1][2] * m[2][1]) * id; inv[1] = (m[0][2] * m[2][1] - m[0][1] * m[2][2]) * id;
        inv[2] = (m[0][1] * m[1][2] - m[0][2] * m[1][1]) * id; inv[3] = (m[1][2] * m[2][0] - m[1][0] * m[2][2]) * id;
        inv[4] = (m[0][0] * m[2][2] - m[0][2] * m[2][0]) * id; inv[5] = (m[0][2] * m[1][0] - m[0][0] * m[1][2]) * id;
        inv[6] = (m[1][0] * m[2][1] - m[1][1] * m[2][0]) * id; inv[7] = (m[0][1] * m[2][0] - m[0][0] * m[2][1]) * id;
        inv[8] = (m[0][0] * m[1][1] - m[0][1] * m[1][0]) * id;
    }
}

// S -= sum_p W_p Vinv_p W_p^T over the camera pairs observing p; rhs += W_p Vinv_p gp.  One thread per point.
__global__ void k_ba_schur(Ba b) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (b.state[4] != 0.0 || p >= b.P) return;
    const double *vi = b.Vinv + p * 9, *g = b.gp + p * 3;
    for (int a = b.pt_off[p]; a < b.pt_off[p + 1]; ++a) {
        const int oa = b.pt_obs[a], ca = b.cam_idx[oa];
        const double *wa = b.Wo + (long long)oa * 18;
        double Y[6][3];
        for (int i = 0; i < 6; ++i)
            for (int j = 0; j < 3; ++j) Y[i][j] = wa[i * 3] * vi[j] + wa[i * 3 + 1] * vi[3 + j] + wa[i * 3 + 2] * vi[6 + j];
        for (int i = 0; i < 6; ++i) atomicAdd(&b.rhs[ca * 6 + i], Y[i][0] * g[0] + Y[i][1] * g[1] + Y[i][2] * g[2]);
        for (int e = b.pt_off[p]; e < b.pt_off[p + 1]; ++e) {
            const int oe = b.pt_obs[e], ce = b.cam_idx[oe];
            const double *we = b.Wo + (long long)oe * 18;
            for (int i = 0; i < 6; ++i)
                for (int j = 0; j < 6; ++j)
                    atomicAdd(&b.S[(long long)(ca * 6 + i) * b.n + ce * 6 + j], -(Y[i][0] * we[j * 3] + Y[i][1] * we[j * 3 + 1] + Y[i][2] * we[j * 3 + 2]));
        }
    }
}

// dense Cholesky S = L L^T in place (lower), then L y = rhs, L^T dc = y.  One CTA.
__global__ void __launch_bounds__(256)
k_ba_solve(Ba b) {
    if (b.state[4] != 0.0) return;
    const int n = b.n, tid = threadIdx.x;
    double *S = b.S;
    __shared__ double diag;
    for (int j = 0; j < n; ++j) {
        if (tid == 0) { const double d = S[(long long)j * n + j]; diag = sqrt(d > 1e-300 ? d : 1e-300); S[(long long)j * n + j] = diag; }
        __syncthreads();
        for (int i = j + 1 + tid; i < n; i += blockDim.x) S[(long long)i * n + j] /= diag;
        __syncthreads();
        const int m = n - j - 1;
        for (int e = tid; e < m * m; e += blockDim.x) {
            const int r = j + 1 + e / m, c = j + 1 + e % m;
            if (c <= r) S[(long long)r * n + c] -= S[(long long)r * n + j] * S[(long long)c * n + j];
        }
        __syncthreads();
    }
    if (tid == 0) {
        double *y = b.dc;
        for (int i = 0; i < n; ++i) {
            double s = b.rhs[i];
            for (int k = 0; k < i; ++k) s -= S[(long long)i * n + k] * y[k];
            y[i] = s / S[(long long)i * n + i];
        }
        for (int i = n - 1; i >= 0; --i) {
            double s = y[i];
            for (int k = i + 1; k < n; ++k) s -= S[(long long)k * n + i] * y[k];
            y[i] = s / S[(long long)i * n + i];
        }
    }
}

// candidate: cams + dc; points + dp with dp = -Vinv (gp + sum_o W_o^T dc_c(o))
__global__ void k_ba_candidate(Ba b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (b.state[4] != 0.0) return;
    if (i < b.C * 6) b.cams_new[i] = b.cams[i] + b.dc[i];
    if (i < b.P) {
        double t[3] = {b.gp[i * 3], b.gp[i * 3 + 1], b.gp[i * 3 + 2]};
        for (int a = b.pt_off[i]; a < b.pt_off[i + 1]; ++a) {
            const int o = b.pt_obs[a], c = b.cam_idx[o];
            const double *w = b.Wo + (long long)o * 18, *d = b.dc + c * 6;
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 6; ++k) t[j] += w[k * 3 + j] * d[k];
        }
        const double *vi = b.Vinv + i * 9;
        for (int j = 0; j < 3; ++j) b.pts_new[i * 3 + j] = b.pts[i * 3 + j] - (vi[j * 3] * t[0] + vi[j * 3 + 1] * t[1] + vi[j * 3 + 2] * t[2]);
    }
}

__global__ void k_ba_eval(Ba b) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (b.state[4] != 0.0) return;
    double cost = 0.0;
    if (o < b.O) {
        double r[2];
        residual(b.cams_new + b.cam_idx[o] * 6, b.pts_new + b.pt_idx[o] * 3, b.feats + o * 5, r);
        cost = r[0] * r[0] + r[1] * r[1];
        if (!(cost == cost)) cost = 1e300;                        // a NaN step is a rejected step
    }
    cost = warp_sum(cost);
    if ((threadIdx.x & 31) == 0 && cost != 0.0) atomicAdd(&b.state[2], cost);
}

__global__ void k_ba_decide(Ba b) {
    if (b.state[4] != 0.0) return;
    if (b.state[6] == 0.0) b.state[7] = b.state[1];
    b.state[6] += 1.0;
    if (b.state[2] < b.state[1]) {
        b.state[5] = 1.0; b.state[3] += 1.0;
        b.state[0] = fmax(b.state[0] * 0.1, 1e-12);
    } else {
        b.state[5] = 0.0;
        b.state[0] = fmin(b.state[0] * 10.0, 1e12);
    }
}

__global__ void k_ba_commit(Ba b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (b.state[4] != 0.0) return;
    if (b.state[5] != 0.0) {
        if (i < b.C * 6) b.cams[i] = b.cams_new[i];
        if (i < b.P * 3) b.pts[i] = b.pts_new[i];
    }
}

__global__ void k_ba_finish_iter(Ba b) {
    if (b.state[4] != 0.0) return;
    if (b.state[5] != 0.0) b.state[1] = b.state[2];
    if (b.state[3] >= (double)b.max_success) b.state[4] = 1.0;
}

}  // namespace trk
}  // namespace opb

extern "C" int op_snavely_residual(const double *points, const double *cams, const double *features, const int64_t *pt_idx, const int64_t *cam_idx,
                                   long long n, double *residual, void *stream) {
    OPB_CHECK_ARG(points && cams && features && residual && n >= 0, "op_snavely_residual: bad argument");
    if (n == 0) return OPB_OK;
    opb::trk::k_snavely_residual<<<(unsigned)opb::ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(points, cams, features, pt_idx, cam_idx, n, residual);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_project_points(const double *xyz, const double *K, const double *RT, long long n, const double *ref2d, double *xy, double *depth,
                                 double *dist_sum, void *stream) {
    OPB_CHECK_ARG(xyz && K && RT && xy && n >= 0 && (!ref2d || dist_sum), "op_project_points: bad argument");
    if (dist_sum) OPB_CHECK_CUDA(cudaMemsetAsync(dist_sum, 0, sizeof(double), (cudaStream_t)stream));
    if (n == 0) return OPB_OK;
    opb::trk::k_project_points<<<(unsigned)opb::ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(xyz, K, RT, n, ref2d, xy, depth, dist_sum);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_triangulate_points(const double *P1, const double *P2, const double *x1, const double *x2, long long n, double *X, void *stream) {
    OPB_CHECK_ARG(P1 && P2 && x1 && x2 && X && n >= 0, "op_triangulate_points: bad argument");
    if (n == 0) return OPB_OK;
    opb::trk::k_triangulate<<<(unsigned)opb::ceil_div(n, 64), 64, 0, (cudaStream_t)stream>>>(P1, P2, x1, x2, n, X);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" size_t op_ba_workspace_bytes(int C, int P, int O) {
    const size_t n = 6 * (size_t)C;
    const size_t d = (size_t)C * 36 + (size_t)P * 9 + (size_t)C * 6 + (size_t)P * 3 + (size_t)O * 18 + (size_t)P * 9 + n * n + n + n + (size_t)C * 6 +
                     (size_t)P * 3 + 8;
    return d * sizeof(double);
}

extern "C" int op_ba_lm(double *cams, double *points, const double *features, const int32_t *cam_idx, const int32_t *pt_idx, const int32_t *pt_off,
                        const int32_t *pt_obs, int C, int P, int O, int num_iterations, int num_success_iterations, double lambda0, void *workspace,
                        size_t workspace_bytes, double *state_out, void *stream) {
    using namespace opb::trk;
    OPB_CHECK_ARG(cams && points && features && cam_idx && pt_idx && pt_off && pt_obs && workspace, "op_ba_lm: null pointer");
    OPB_CHECK_ARG(C > 0 && C <= 64 && P > 0 && O > 0 && num_iterations >= 0 && lambda0 > 0.0, "op_ba_lm: bad size (1..64 cameras)");
    OPB_CHECK_ARG(workspace_bytes >= op_ba_workspace_bytes(C, P, O), "op_ba_lm: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    Ba b{};
    b.C = C; b.P = P; b.O = O; b.n = 6 * C; b.cams = cams; b.pts = points; b.feats = features; b.cam_idx = cam_idx; b.pt_idx = pt_idx;
    b.pt_off = pt_off; b.pt_obs = pt_obs; b.max_success = num_success_iterations;
    double *w = (double *)workspace;
    b.U = w; w += (size_t)C * 36; b.V = w; w += (size_t)P * 9; b.gc = w; w += (size_t)C * 6; b.gp = w; w += (size_t)P * 3;
    b.Wo = w; w += (size_t)O * 18; b.Vinv = w; w += (size_t)P * 9; b.S = w; w += (size_t)b.n * b.n; b.rhs = w; w += b.n; b.dc = w; w += b.n;
    b.cams_new = w; w += (size_t)C * 6; b.pts_new = w; w += (size_t)P * 3; b.state = w;
    const double init[8] = {lambda0, 0, 0, 0, 0, 0, 0, 0};
    OPB_CHECK_CUDA(cudaMemcpyAsync(b.state, init, sizeof(init), cudaMemcpyHostToDevice, st));
    int big = C * 36;
    if (P * 9 > big) big = P * 9;
    if (b.n * b.n > big) big = b.n * b.n;
    const int gz = opb::ceil_div(big, 256), go = opb::ceil_div(O, 128), gp = opb::ceil_div(P > C * 6 ? P : C * 6, 128), g3 = opb::ceil_div(P * 3 > C * 6 ? P * 3 : C * 6, 128);
    for (int it = 0; it < num_iterations; ++it) {
        k_ba_zero<<<gz, 256, 0, st>>>(b);
        k_ba_linearize<<<go, 128, 0, st>>>(b);
        k_ba_blocks<<<gp, 128, 0, st>>>(b);
        k_ba_schur<<<opb::ceil_div(P, 128), 128, 0, st>>>(b);
        k_ba_solve<<<1, 256, 0, st>>>(b);
        k_ba_candidate<<<gp, 128, 0, st>>>(b);
        k_ba_eval<<<go, 128, 0, st>>>(b);
        k_ba_decide<<<1, 1, 0, st>>>(b);
        k_ba_commit<<<g3, 128, 0, st>>>(b);
        k_ba_finish_iter<<<1, 1, 0, st>>>(b);
    }
    OPB_CHECK_LAUNCH();
    if (state_out) OPB_CHECK_CUDA(cudaMemcpyAsync(state_out, b.state, 8 * sizeof(double), cudaMemcpyDeviceToDevice, st));
    return OPB_OK;
}
