// EPnP stages shared by the per-hypothesis (thread-sequential, 5 points) and the refit
// (block-parallel, all inliers) paths of pnp.cu.  See pnp_math.cuh for provenance.
#pragma once
#include "pnp_math.cuh"

namespace opb {
namespace pnp {

// L (6x10) from the four smallest singular vectors (rows 11,10,9,8 of ut)
__device__ inline void compute_L_6x10(const double *ut, double *L) {
    const double *v[4] = {ut + 12 * 11, ut + 12 * 10, ut + 12 * 9, ut + 12 * 8};
    double dv[4][6][3];
    for (int i = 0; i < 4; ++i) {
        int a = 0, b = 1;
        for (int j = 0; j < 6; ++j) {
            for (int k = 0; k < 3; ++k) dv[i][j][k] = v[i][3 * a + k] - v[i][3 * b + k];
            if (++b > 3) { ++a; b = a + 1; }
        }
    }
    for (int i = 0; i < 6; ++i) {
        double *row = L + 10 * i;
        row[0] = dot3(dv[0][i], dv[0][i]);
        row[1] = 2.0 * dot3(dv[0][i], dv[1][i]);
        row[2] = dot3(dv[1][i], dv[1][i]);
        row[3] = 2.0 * dot3(dv[0][i], dv[2][i]);
        row[4] = 2.0 * dot3(dv[1][i], dv[2][i]);
        row[5] = dot3(dv[2][i], dv[2][i]);
        row[6] = 2.0 * dot3(dv[0][i], dv[3][i]);
        row[7] = 2.0 * dot3(dv[1][i], dv[3][i]);
        row[8] = 2.0 * dot3(dv[2][i], dv[3][i]);
        row[9] = dot3(dv[3][i], dv[3][i]);
    }
}

__device__ inline void compute_rho(const double cws[4][3], double *rho) {
    rho[0] = dist2(cws[0], cws[1]); rho[1] = dist2(cws[0], cws[2]); rho[2] = dist2(cws[0], cws[3]);
    rho[3] = dist2(cws[1], cws[2]); rho[4] = dist2(cws[1], cws[3]); rho[5] = dist2(cws[2], cws[3]);
}

__device__ inline void betas_approx_1(const double *L, const double *rho, double *betas) {
    double A[24], b4[4];
    for (int i = 0; i < 6; ++i) { A[4 * i] = L[10 * i]; A[4 * i + 1] = L[10 * i + 1]; A[4 * i + 2] = L[10 * i + 3]; A[4 * i + 3] = L[10 * i + 6]; }
    svd_lstsq6<4>(A, rho, b4);
    if (b4[0] < 0) { betas[0] = sqrt(-b4[0]); betas[1] = -b4[1] / betas[0]; betas[2] = -b4[2] / betas[0]; betas[3] = -b4[3] / betas[0]; }
    else { betas[0] = sqrt(b4[0]); betas[1] = b4[1] / betas[0]; betas[2] = b4[2] / betas[0]; betas[3] = b4[3] / betas[0]; }
}
__device__ inline void betas_approx_2(const double *L, const double *rho, double *betas) {
    double A[18], b3[3];
    for (int i = 0; i < 6; ++i) { A[3 * i] = L[10 * i]; A[3 * i + 1] = L[10 * i + 1]; A[3 * i + 2] = L[10 * i + 2]; }
    svd_lstsq6<3>(A, rho, b3);
    if (b3[0] < 0) { betas[0] = sqrt(-b3[0]); betas[1] = (b3[2] < 0) ? sqrt(-b3[2]) : 0.0; }
    else { betas[0] = sqrt(b3[0]); betas[1] = (b3[2] > 0) ? sqrt(b3[2]) : 0.0; }
    if (b3[1] < 0) betas[0] = -betas[0];
    betas[2] = 0.0; betas[3] = 0.0;
}
__device__ inline void betas_approx_3(const double *L, const double *rho, double *betas) {
    double A[30], b5[5];
    for (int i = 0; i < 6; ++i) for (int k = 0; k < 5; ++k) A[5 * i + k] = L[10 * i + k];
    svd_lstsq6<5>(A, rho, b5);
    if (b5[0] < 0) { betas[0] = sqrt(-b5[0]); betas[1] = (b5[2] < 0) ? sqrt(-b5[2]) : 0.0; }
    else { betas[0] = sqrt(b5[0]); betas[1] = (b5[2] > 0) ? sqrt(b5[2]) : 0.0; }
    if (b5[1] < 0) betas[0] = -betas[0];
    betas[2] = b5[3] / betas[0];
    betas[3] = 0.0;
}

__device__ inline void gauss_newton(const double *L, const double *rho, double *bt) {
    for (int it = 0; it < 5; ++it) {
        double A[24], b[6], x[4];
        for (int i = 0; i < 6; ++i) {
            const double *r = L + 10 * i;
            double *ra = A + 4 * i;
            ra[0] = 2 * r[0] * bt[0] + r[1] * bt[1] + r[3] * bt[2] + r[6] * bt[3];
            ra[1] = r[1] * bt[0] + 2 * r[2] * bt[1] + r[4] * bt[2] + r[7] * bt[3];
            ra[2] = r[3] * bt[0] + r[4] * bt[1] + 2 * r[5] * bt[2] + r[8] * bt[3];
            ra[3] = r[6] * bt[0] + r[7] * bt[1] + r[8] * bt[2] + 2 * r[9] * bt[3];
            b[i] = rho[i] - (r[0] * bt[0] * bt[0] + r[1] * bt[0] * bt[1] + r[2] * bt[1] * bt[1] + r[3] * bt[0] * bt[2] +
                             r[4] * bt[1] * bt[2] + r[5] * bt[2] * bt[2] + r[6] * bt[0] * bt[3] + r[7] * bt[1] * bt[3] +
                             r[8] * bt[2] * bt[3] + r[9] * bt[3] * bt[3]);
        }
        qr_solve_6x4(A, b, x);
        for (int i = 0; i < 4; ++i) bt[i] += x[i];
    }
}

// the three (beta initialisation + Gauss-Newton) candidates; betas is [3][4]
__device__ inline void solve_betas(const double *ut, const double cws[4][3], double betas[3][4]) {
    double L[60], rho[6];
    compute_L_6x10(ut, L);
    compute_rho(cws, rho);
    betas_approx_1(L, rho, betas[0]); gauss_newton(L, rho, betas[0]);
    betas_approx_2(L, rho, betas[1]); gauss_newton(L, rho, betas[1]);
    betas_approx_3(L, rho, betas[2]); gauss_newton(L, rho, betas[2]);
}

__device__ inline void compute_ccs(const double *ut, const double *betas, double ccs[4][3]) {
    for (int i = 0; i < 4; ++i) ccs[i][0] = ccs[i][1] = ccs[i][2] = 0;
    for (int i = 0; i < 4; ++i) {
        const double *v = ut + 12 * (11 - i);
        for (int j = 0; j < 4; ++j) for (int k = 0; k < 3; ++k) ccs[j][k] += betas[i] * v[3 * j + k];
    }
}

// control points from the centroid and the (symmetric) 3x3 scatter c = PW0^T PW0
__device__ inline void control_points_from_scatter(double *c /*3x3, destroyed*/, int n, double cws[4][3]) {
    double dc[3];
    jacobi_svd<3, 3, false>(c, dc, nullptr);
    for (int i = 1; i < 4; ++i) {
        const double k = sqrt(dc[i - 1] / n);
        for (int j = 0; j < 3; ++j) cws[i][j] = cws[0][j] + k * c[3 * (i - 1) + j];
    }
}

// Procrustes step: R = U V^T from abt (3x3, row-major: abt[3*j+k] = sum (pc_j - pc0_j)(pw_k - pw0_k))
__device__ inline void rotation_from_abt(const double *abt, const double *pc0, const double *pw0, double *R, double *t) {
    double At[9], d[3], Vt[9];
    for (int i = 0; i < 3; ++i) for (int k = 0; k < 3; ++k) At[3 * i + k] = abt[3 * k + i];
    jacobi_svd<3, 3, true>(At, d, Vt);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j)
        R[3 * i + j] = At[0 + i] * Vt[0 + j] + At[3 + i] * Vt[3 + j] + At[6 + i] * Vt[6 + j];
    const double det = R[0] * R[4] * R[8] + R[1] * R[5] * R[6] + R[2] * R[3] * R[7] - R[2] * R[4] * R[6] - R[1] * R[3] * R[8] - R[0] * R[5] * R[7];
    if (det < 0) { R[6] = -R[6]; R[7] = -R[7]; R[8] = -R[8]; }
    for (int j = 0; j < 3; ++j) t[j] = pc0[j] - dot3(R + 3 * j, pw0);
}

// Thread-sequential EPnP for a minimal sample (NP points), identity camera (normalised image coords).
template <int NP>
__device__ void epnp_small(const double *pws, const double *us, double *Rout, double *tout) {
    double cws[4][3], alphas[NP][4];
    cws[0][0] = cws[0][1] = cws[0][2] = 0;
    for (int i = 0; i < NP; ++i) for (int j = 0; j < 3; ++j) cws[0][j] += pws[3 * i + j];
    for (int j = 0; j < 3; ++j) cws[0][j] /= NP;
    double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < NP; ++i) {
        double d[3];
        for (int j = 0; j < 3; ++j) d[j] = pws[3 * i + j] - cws[0][j];
        for (int r = 0; r < 3; ++r) for (int s = r; s < 3; ++s) c[3 * r + s] += d[r] * d[s];
    }
    c[3] = c[1]; c[6] = c[2]; c[7] = c[5];
    control_points_from_scatter(c, NP, cws);
    double cc[9], ci[9];
    for (int i = 0; i < 3; ++i) for (int j = 1; j < 4; ++j) cc[3 * i + j - 1] = cws[j][i] - cws[0][i];
    inv3(cc, ci);
    for (int i = 0; i < NP; ++i) {
        const double *pi = pws + 3 * i;
        double *a = alphas[i];
        for (int j = 0; j < 3; ++j)
            a[1 + j] = ci[3 * j] * (pi[0] - cws[0][0]) + ci[3 * j + 1] * (pi[1] - cws[0][1]) + ci[3 * j + 2] * (pi[2] - cws[0][2]);
        a[0] = 1.0 - a[1] - a[2] - a[3];
    }
    double mtm[144];
    for (int i = 0; i < 144; ++i) mtm[i] = 0;
    for (int i = 0; i < NP; ++i) {
        const double *as = alphas[i];
        const double u = us[2 * i], v = us[2 * i + 1];
        double m1[12], m2[12];
        for (int k = 0; k < 4; ++k) {
            m1[3 * k] = as[k] * 1.0; m1[3 * k + 1] = 0.0; m1[3 * k + 2] = as[k] * (0.0 - u);
            m2[3 * k] = 0.0; m2[3 * k + 1] = as[k] * 1.0; m2[3 * k + 2] = as[k] * (0.0 - v);
        }
        for (int r = 0; r < 12; ++r) for (int s = r; s < 12; ++s) mtm[12 * r + s] += m1[r] * m1[s] + m2[r] * m2[s];
    }
    for (int r = 0; r < 12; ++r) for (int s = 0; s < r; ++s) mtm[12 * r + s] = mtm[12 * s + r];
    double dsv[12];
    jacobi_svd<12, 12, false>(mtm, dsv, nullptr);     // rows of mtm are now ut
    double betas[3][4];
    solve_betas(mtm, cws, betas);
    double best = 0; int have = 0;
    for (int cand = 0; cand < 3; ++cand) {
        double ccs[4][3], pcs[NP][3], R[9], t[3];
        compute_ccs(mtm, betas[cand], ccs);
        for (int i = 0; i < NP; ++i)
            for (int j = 0; j < 3; ++j)
                pcs[i][j] = alphas[i][0] * ccs[0][j] + alphas[i][1] * ccs[1][j] + alphas[i][2] * ccs[2][j] + alphas[i][3] * ccs[3][j];
        if (pcs[0][2] < 0.0)
            for (int i = 0; i < NP; ++i) for (int j = 0; j < 3; ++j) pcs[i][j] = -pcs[i][j];
        double pc0[3] = {0, 0, 0}, pw0[3] = {0, 0, 0};
        for (int i = 0; i < NP; ++i) for (int j = 0; j < 3; ++j) { pc0[j] += pcs[i][j]; pw0[j] += pws[3 * i + j]; }
        for (int j = 0; j < 3; ++j) { pc0[j] /= NP; pw0[j] /= NP; }
        double abt[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < NP; ++i) {
            const double *pw = pws + 3 * i;
            for (int j = 0; j < 3; ++j) {
                abt[3 * j] += (pcs[i][j] - pc0[j]) * (pw[0] - pw0[0]);
                abt[3 * j + 1] += (pcs[i][j] - pc0[j]) * (pw[1] - pw0[1]);
                abt[3 * j + 2] += (pcs[i][j] - pc0[j]) * (pw[2] - pw0[2]);
            }
        }
        rotation_from_abt(abt, pc0, pw0, R, t);
        double sum2 = 0;
        for (int i = 0; i < NP; ++i) {
            const double *pw = pws + 3 * i;
            const double Xc = dot3(R, pw) + t[0], Yc = dot3(R + 3, pw) + t[1], inv_Zc = 1.0 / (dot3(R + 6, pw) + t[2]);
            const double ue = 0.0 + 1.0 * Xc * inv_Zc, ve = 0.0 + 1.0 * Yc * inv_Zc;
            const double u = us[2 * i], v = us[2 * i + 1];
            sum2 += sqrt((u - ue) * (u - ue) + (v - ve) * (v - ve));
        }
        const double rep = sum2 / NP;
        // best-of-three: candidate 2 replaces 1 iff strictly smaller, candidate 3 likewise
        if (!have || rep < best) {
            best = rep; have = 1;
            for (int i = 0; i < 9; ++i) Rout[i] = R[i];
            for (int i = 0; i < 3; ++i) tout[i] = t[i];
        }
    }
}

}  // namespace pnp
}  // namespace opb
