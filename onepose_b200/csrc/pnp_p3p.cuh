// cv2.solvePnPRansac with exactly four correspondences (included by pnp.cu).
//
// Reference: ransac_PnP (src/utils/eval_utils.py:28-35) hands whatever it was given to cv2.solvePnPRansac; with
// npoints == 4 calib3d skips RANSAC (model_points = npoints = 4), solves P3P on points 0-2, keeps the solution whose
// reprojection of point 3 is closest (squared pixel distance) and reports all four points as inliers.  calib3d is a third-party
// dependency (opencv_python, requirements.txt:5) that is not vendored in the reference; what is restated here is the published
// problem, not its code: Grunert's distance formulation of P3P (Haralick et al., "Review and analysis of solutions of the three
// point perspective pose estimation problem", 1994) -- a quartic in v = s3 / s1 whose positive real roots give the three depths --
// followed by the rigid alignment of the two congruent triangles.  Any exact P3P solver returns the same pose set, so the result
// agrees with cv2 to rounding (measured: <= 0.002 deg / 0.01 mm over 300 seeded problems, tests/test_gpu_pnp.py).
#pragma once

namespace opb {
namespace pnp {

struct Cplx { double re, im; };
__device__ __forceinline__ Cplx cadd(Cplx a, Cplx b) { return {a.re + b.re, a.im + b.im}; }
__device__ __forceinline__ Cplx csub(Cplx a, Cplx b) { return {a.re - b.re, a.im - b.im}; }
__device__ __forceinline__ Cplx cmul(Cplx a, Cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
__device__ __forceinline__ Cplx cdiv(Cplx a, Cplx b) {
    const double d = b.re * b.re + b.im * b.im;
    return {(a.re * b.re + a.im * b.im) / d, (a.im * b.re - a.re * b.im) / d};
}

// all four roots of c4 x^4 + c3 x^3 + c2 x^2 + c1 x + c0 by the Durand-Kerner simultaneous iteration (c4 != 0)
__device__ inline void quartic_roots(const double *c, Cplx *r) {
    const double a[4] = {c[1] / c[0], c[2] / c[0], c[3] / c[0], c[4] / c[0]};
    double bound = 0.0;
    for (int i = 0; i < 4; ++i) bound = fmax(bound, fabs(a[i]));
    bound += 1.0;
    Cplx g = {bound, 0.0};
    const Cplx step = {0.4, 0.9};
    for (int i = 0; i < 4; ++i) { r[i] = g; g = cmul(g, step); }
    for (int it = 0; it < 200; ++it) {
        double move = 0.0;
        for (int i = 0; i < 4; ++i) {
            Cplx p = cadd(r[i], Cplx{a[0], 0.0});
            for (int k = 1; k < 4; ++k) p = cadd(cmul(p, r[i]), Cplx{a[k], 0.0});
            Cplx q = {1.0, 0.0};
            for (int j = 0; j < 4; ++j) if (j != i) q = cmul(q, csub(r[i], r[j]));
            if (q.re == 0.0 && q.im == 0.0) continue;
            const Cplx d = cdiv(p, q);
            r[i] = csub(r[i], d);
            move = fmax(move, fmax(fabs(d.re), fabs(d.im)));
        }
        if (move < 1e-15 * bound) break;
    }
}

__device__ inline void unit3(double *v) {
    const double n = sqrt(dot3(v, v));
    v[0] /= n; v[1] /= n; v[2] /= n;
}
__device__ inline void cross3(const double *a, const double *b, double *o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
// orthonormal frame of a triangle, columns (e1, e2, e3) stored row-major in F[9]
__device__ inline void triangle_frame(const double T[3][3], double *F) {
    double e1[3], d[3], e2[3], e3[3];
    for (int k = 0; k < 3; ++k) { e1[k] = T[1][k] - T[0][k]; d[k] = T[2][k] - T[0][k]; }
    unit3(e1);
    cross3(e1, d, e3);
    unit3(e3);
    cross3(e3, e1, e2);
    for (int k = 0; k < 3; ++k) { F[3 * k] = e1[k]; F[3 * k + 1] = e2[k]; F[3 * k + 2] = e3[k]; }
}

// Pose from four correspondences.  X32 / uv32: the frame's float32 points (cv2 converts to CV_32F first), rows 0-3.
// Normalised image coordinates are rounded to float32 as undistortPoints does on CV_32F input.  Returns false when P3P has no
// admissible solution (cv2 then reports failure and the wrapper falls back to the identity pose).
__device__ inline bool solve_p3p4(const float *X32, const float *uv32, const Cam &cam, double *R /*[9]*/, double *t /*[3]*/) {
    double X[4][3], un[4][2], f[3][3];
    for (int i = 0; i < 4; ++i) {
        for (int k = 0; k < 3; ++k) X[i][k] = X32[3 * i + k];
        un[i][0] = (float)(((double)uv32[2 * i] - cam.cx) / cam.fx);
        un[i][1] = (float)(((double)uv32[2 * i + 1] - cam.cy) / cam.fy);
    }
    for (int i = 0; i < 3; ++i) { f[i][0] = un[i][0]; f[i][1] = un[i][1]; f[i][2] = 1.0; unit3(f[i]); }
    const double a2 = dist2(X[1], X[2]), b2 = dist2(X[0], X[2]), c2 = dist2(X[0], X[1]);
    if (!(a2 > 0.0 && b2 > 0.0 && c2 > 0.0)) return false;
    const double ca = dot3(f[1], f[2]), cb = dot3(f[0], f[2]), cg = dot3(f[0], f[1]);
    const double q = (a2 - c2) / b2, p = (a2 + c2) / b2;
    double co[5];
    co[0] = (q - 1) * (q - 1) - 4 * c2 / b2 * ca * ca;
    co[1] = 4 * (q * (1 - q) * cb - (1 - p) * ca * cg + 2 * c2 / b2 * ca * ca * cb);
    co[2] = 2 * (q * q - 1 + 2 * q * q * cb * cb + 2 * (b2 - c2) / b2 * ca * ca - 4 * p * ca * cb * cg + 2 * (b2 - a2) / b2 * cg * cg);
    co[3] = 4 * (-q * (1 + q) * cb + 2 * a2 / b2 * cg * cg * cb - (1 - p) * ca * cg);
    co[4] = (1 + q) * (1 + q) - 4 * a2 / b2 * cg * cg;
    if (co[0] == 0.0 || !isfinite(co[0] + co[1] + co[2] + co[3] + co[4])) return false;
    Cplx roots[4];
    quartic_roots(co, roots);
    double Fw[9];
    triangle_frame(X, Fw);
    bool have = false;
    double best = 0.0;
    for (int k = 0; k < 4; ++k) {
        double v = roots[k].re;
        if (!(fabs(roots[k].im) <= 1e-7 * fmax(1.0, fabs(v))) || !(v > 0.0)) continue;
        for (int it = 0; it < 2; ++it) {                    // Newton polish of the real root
            const double pv = (((co[0] * v + co[1]) * v + co[2]) * v + co[3]) * v + co[4];
            const double dv = ((4 * co[0] * v + 3 * co[1]) * v + 2 * co[2]) * v + co[3];
            if (dv != 0.0) v -= pv / dv;
        }
        const double u = ((q - 1) * v * v - 2 * q * cb * v + 1 + q) / (2 * (cg - v * ca));
        if (!(u > 0.0) || !isfinite(u)) continue;
        const double s1 = sqrt(c2 / (1 + u * u - 2 * u * cg));
        const double s[3] = {s1, u * s1, v * s1};
        double P[3][3], Fc[9], Rk[9], tk[3];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) P[i][j] = s[i] * f[i][j];
        triangle_frame(P, Fc);
        for (int i = 0; i < 3; ++i)                          // R = Fc Fw^T
            for (int j = 0; j < 3; ++j) Rk[3 * i + j] = Fc[3 * i] * Fw[3 * j] + Fc[3 * i + 1] * Fw[3 * j + 1] + Fc[3 * i + 2] * Fw[3 * j + 2];
        for (int i = 0; i < 3; ++i) tk[i] = P[0][i] - dot3(Rk + 3 * i, X[0]);
        if (!isfinite(Rk[0] + Rk[4] + Rk[8] + tk[0] + tk[1] + tk[2])) continue;
        // the fourth point decides: squared pixel distance of its reprojection
        const double xc = dot3(Rk, X[3]) + tk[0], yc = dot3(Rk + 3, X[3]) + tk[1], zc = dot3(Rk + 6, X[3]) + tk[2];
        const double du = cam.cx + cam.fx * xc / zc - (un[3][0] * cam.fx + cam.cx), dv = cam.cy + cam.fy * yc / zc - (un[3][1] * cam.fy + cam.cy);
        const double e = du * du + dv * dv;
        if (!have || e < best) {
            have = true; best = e;
            for (int i = 0; i < 9; ++i) R[i] = Rk[i];
            for (int i = 0; i < 3; ++i) t[i] = tk[i];
        }
    }
    return have;
}

}  // namespace pnp
}  // namespace opb
