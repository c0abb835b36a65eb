// Device-side float64 building blocks of the EPnP solver used by pnp.cu.
//
// The reference calls cv2.solvePnPRansac (src/utils/eval_utils.py:28-29); OpenCV's calib3d is a
// third-party dependency that is not vendored in the reference tree, so these routines implement the
// published algorithm (Lepetit, Moreno-Noguer, Fua: "EPnP", IJCV 2009) in the formulation calib3d uses:
// one-sided Jacobi SVD for the small dense decompositions, SVD back-substitution for the three
// beta initialisations, Householder QR for the 6x4 Gauss-Newton step.  Compiled with -fmad=false so
// a*b+c is never contracted (OpenCV's x86 build does not use FMA), which keeps hypothesis poses
// bit-close to the library on well-posed samples.
#pragma once
#include <float.h>
#include <math.h>

namespace opb {
namespace pnp {

__device__ __forceinline__ double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
__device__ __forceinline__ double dist2(const double *a, const double *b) {
    return (a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]) + (a[2] - b[2]) * (a[2] - b[2]);
}

// One-sided (Hestenes) Jacobi SVD.  At holds n rows of length m (row i = column i of A) and is
// overwritten by the normalised left singular vectors; W gets the singular values (descending);
// Vt (n x n, optional) the right singular vectors as rows.
template <int M, int N, bool WITH_V>
__device__ void jacobi_svd(double *At, double *Wout, double *Vt) {
    const double eps = DBL_EPSILON * 10, minval = DBL_MIN;
    double W[N];
    const int max_iter = M > 30 ? M : 30;
    for (int i = 0; i < N; ++i) {
        double sd = 0;
        for (int k = 0; k < M; ++k) sd += At[i * M + k] * At[i * M + k];
        W[i] = sd;
        if (WITH_V) for (int k = 0; k < N; ++k) Vt[i * N + k] = (i == k) ? 1.0 : 0.0;
    }
    for (int iter = 0; iter < max_iter; ++iter) {
        bool changed = false;
        for (int i = 0; i < N - 1; ++i)
            for (int j = i + 1; j < N; ++j) {
                double *Ai = At + i * M, *Aj = At + j * M;
                double a = W[i], p = 0, b = W[j];
                for (int k = 0; k < M; ++k) p += Ai[k] * Aj[k];
                if (fabs(p) <= eps * sqrt(a * b)) continue;
                p *= 2;
                const double beta = a - b, gamma = hypot(p, beta);
                double c, s;
                if (beta < 0) {
                    const double delta = (gamma - beta) * 0.5;
                    s = sqrt(delta / gamma);
                    c = p / (gamma * s * 2);
                } else {
                    c = sqrt((gamma + beta) / (gamma * 2));
                    s = p / (gamma * c * 2);
                }
                a = b = 0;
                for (int k = 0; k < M; ++k) {
                    const double t0 = c * Ai[k] + s * Aj[k];
                    const double t1 = -s * Ai[k] + c * Aj[k];
                    Ai[k] = t0; Aj[k] = t1;
                    a += t0 * t0; b += t1 * t1;
                }
                W[i] = a; W[j] = b;
                changed = true;
                if (WITH_V) {
                    double *Vi = Vt + i * N, *Vj = Vt + j * N;
                    for (int k = 0; k < N; ++k) {
                        const double t0 = c * Vi[k] + s * Vj[k];
                        const double t1 = -s * Vi[k] + c * Vj[k];
                        Vi[k] = t0; Vj[k] = t1;
                    }
                }
            }
        if (!changed) break;
    }
    for (int i = 0; i < N; ++i) {
        double sd = 0;
        for (int k = 0; k < M; ++k) sd += At[i * M + k] * At[i * M + k];
        W[i] = sqrt(sd);
    }
    for (int i = 0; i < N - 1; ++i) {
        int j = i;
        for (int k = i + 1; k < N; ++k) if (W[j] < W[k]) j = k;
        if (i != j) {
            double t = W[i]; W[i] = W[j]; W[j] = t;
            for (int k = 0; k < M; ++k) { t = At[i * M + k]; At[i * M + k] = At[j * M + k]; At[j * M + k] = t; }
            if (WITH_V) for (int k = 0; k < N; ++k) { t = Vt[i * N + k]; Vt[i * N + k] = Vt[j * N + k]; Vt[j * N + k] = t; }
        }
    }
    for (int i = 0; i < N; ++i) {
        Wout[i] = W[i];
        const double s = W[i] > minval ? 1.0 / W[i] : 0.0;
        for (int k = 0; k < M; ++k) At[i * M + k] *= s;
    }
}

// x = argmin |A x - b| for a 6 x N system (row-major A) by SVD back-substitution
template <int N>
__device__ void svd_lstsq6(const double *A, const double *b, double *x) {
    double At[N * 6], W[N], Vt[N * N];
    for (int i = 0; i < N; ++i) for (int k = 0; k < 6; ++k) At[i * 6 + k] = A[k * N + i];
    jacobi_svd<6, N, true>(At, W, Vt);
    double thr = 0;
    for (int i = 0; i < N; ++i) thr += W[i];
    thr *= DBL_EPSILON * 2;
    for (int k = 0; k < N; ++k) x[k] = 0;
    for (int i = 0; i < N; ++i) {
        if (W[i] <= thr) continue;
        double ub = 0;
        for (int k = 0; k < 6; ++k) ub += At[i * 6 + k] * b[k];
        ub /= W[i];
        for (int k = 0; k < N; ++k) x[k] += ub * Vt[i * N + k];
    }
}

__device__ inline void qr_solve_6x4(double *A, double *b, double *X) {
    const int nr = 6, nc = 4;
    double A1[4], A2[4];
    for (int k = 0; k < nc; ++k) {
        double eta = 0;
        for (int i = k; i < nr; ++i) { const double e = fabs(A[i * nc + k]); if (e > eta) eta = e; }
        if (eta == 0) { for (int i = 0; i < nc; ++i) X[i] = 0; return; }
        double sum2 = 0;
        const double inv_eta = 1.0 / eta;
        for (int i = k; i < nr; ++i) { A[i * nc + k] *= inv_eta; sum2 += A[i * nc + k] * A[i * nc + k]; }
        double sigma = sqrt(sum2);
        if (A[k * nc + k] < 0) sigma = -sigma;
        A[k * nc + k] += sigma;
        A1[k] = sigma * A[k * nc + k];
        A2[k] = -eta * sigma;
        for (int j = k + 1; j < nc; ++j) {
            double sum = 0;
            for (int i = k; i < nr; ++i) sum += A[i * nc + k] * A[i * nc + j];
            const double tau = sum / A1[k];
            for (int i = k; i < nr; ++i) A[i * nc + j] -= tau * A[i * nc + k];
        }
    }
    for (int j = 0; j < nc; ++j) {
        double tau = 0;
        for (int i = j; i < nr; ++i) tau += A[i * nc + j] * b[i];
        tau /= A1[j];
        for (int i = j; i < nr; ++i) b[i] -= tau * A[i * nc + j];
    }
    X[nc - 1] = b[nc - 1] / A2[nc - 1];
    for (int i = nc - 2; i >= 0; --i) {
        double sum = 0;
        for (int j = i + 1; j < nc; ++j) sum += A[i * nc + j] * X[j];
        X[i] = (b[i] - sum) / A2[i];
    }
}

__device__ inline void inv3(const double *m, double *o) {
    const double det = m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
    const double id = 1.0 / det;
    o[0] = (m[4] * m[8] - m[5] * m[7]) * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = (m[5] * m[6] - m[3] * m[8]) * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = (m[3] * m[7] - m[4] * m[6]) * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

// rotation matrix <-> axis-angle, as cv::Rodrigues
__device__ inline void rodrigues_to_vec(const double *R, double *r) {
    double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
    const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
    double c = (R[0] + R[4] + R[8] - 1) * 0.5;
    c = c > 1. ? 1. : c < -1. ? -1. : c;
    double theta = acos(c);
    if (s < 1e-5) {
        if (c > 0) { r[0] = r[1] = r[2] = 0; return; }
        double t;
        t = (R[0] + 1) * 0.5; rx = sqrt(t > 0 ? t : 0);
        t = (R[4] + 1) * 0.5; ry = sqrt(t > 0 ? t : 0) * (R[1] < 0 ? -1. : 1.);
        t = (R[8] + 1) * 0.5; rz = sqrt(t > 0 ? t : 0) * (R[2] < 0 ? -1. : 1.);
        if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0) != (ry * rz > 0)) rz = -rz;
        theta /= sqrt(rx * rx + ry * ry + rz * rz);
        r[0] = rx * theta; r[1] = ry * theta; r[2] = rz * theta;
    } else {
        double vth = 1 / (2 * s);
        vth *= theta;
        r[0] = rx * vth; r[1] = ry * vth; r[2] = rz * vth;
    }
}
__device__ inline void rodrigues_to_mat(const double *r, double *R) {
    const double theta = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    if (theta < DBL_EPSILON) { for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0 : 0.0; return; }
    const double c = cos(theta), s = sin(theta), c1 = 1. - c, it = 1. / theta;
    const double x = r[0] * it, y = r[1] * it, z = r[2] * it;
    R[0] = c + c1 * x * x; R[1] = c1 * x * y - s * z; R[2] = c1 * x * z + s * y;
    R[3] = c1 * x * y + s * z; R[4] = c + c1 * y * y; R[5] = c1 * y * z - s * x;
    R[6] = c1 * x * z - s * y; R[7] = c1 * y * z + s * x; R[8] = c + c1 * z * z;
}

}  // namespace pnp
}  // namespace opb
