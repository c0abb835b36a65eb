// Second stage of the batched PnP-RANSAC (included by pnp.cu): sequential-equivalent selection,
// inlier mask of the winner, EPnP refit on all inliers, pose output; plus the C entry points.
#pragma once

namespace opb {
namespace pnp {

struct RefitShared {
    double scratch[8 * 78];
    double red[78];
    double cws[4][3], ci[9], ut[144], betas[3][4], ccs[4][3], R[9], t[3], pw0[3];
    double best_R[9], best_t[3], best_rep;
    double hyp_rt[6];
    int best_hyp, iters, n_in, have_best;
    int warp_tot[8], base;
};

__global__ void __launch_bounds__(kRefitThreads)
k_pnp_select_refit(const float *__restrict__ X32, const float *__restrict__ uv32, const int32_t *__restrict__ count, int stride,
                   const double *__restrict__ Kmat, const int32_t *__restrict__ hyp_counts, const double *__restrict__ hyp_pose,
                   int H, float thr2, double confidence, int max_iters, double scale,
                   double *__restrict__ alphas_ws, int32_t *__restrict__ inl_ws,
                   double *__restrict__ pose, uint8_t *__restrict__ inlier_mask, int32_t *__restrict__ n_inliers,
                   int32_t *__restrict__ iters_run) {
    __shared__ RefitShared S;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = count[f];
    const Cam cam = load_cam(Kmat + (long long)f * 9);
    const float *X = X32 + (long long)f * stride * 3, *uv = uv32 + (long long)f * stride * 2;
    uint8_t *mask = inlier_mask + (long long)f * stride;
    double *alphas = alphas_ws + (long long)f * stride * 4;
    int32_t *inl = inl_ws + (long long)f * stride;
    double *out = pose + (long long)f * 12;

    for (int i = tid; i < stride; i += blockDim.x) mask[i] = 0;
    auto write_identity = [&]() {
        if (tid == 0) {
            for (int i = 0; i < 12; ++i) out[i] = (i % 5 == 0) ? 1.0 : 0.0;      // np.eye(4)[:3], eval_utils.py:42
            n_inliers[f] = 0; iters_run[f] = 0;
        }
    };
    // cv2 raises for n < 4 (-> identity pose, no inliers, eval_utils.py:40-42)
    if (n < 4) { write_identity(); return; }
    if (n == 4) {
        // cv2's npoints == 4 branch: no RANSAC, P3P on points 0-2 disambiguated by point 3, all four points reported as inliers
        // (pnp_p3p.cuh); no admissible solution = cv2 failure = the wrapper's identity fallback
        __syncthreads();                              // the mask is cleared
        if (tid == 0) {
            double R4[9], t4[3], rv[3];
            if (solve_p3p4(X, uv, cam, R4, t4)) {
                rodrigues_to_vec(R4, rv);
                rodrigues_to_mat(rv, R4);             // the wrapper's cv2.Rodrigues(rvec), eval_utils.py:31
                for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) out[r * 4 + c] = R4[3 * r + c]; out[r * 4 + 3] = t4[r] / scale; }
                n_inliers[f] = 4; iters_run[f] = 0;
                for (int i = 0; i < 4; ++i) mask[i] = 1;
            } else {
                for (int i = 0; i < 12; ++i) out[i] = (i % 5 == 0) ? 1.0 : 0.0;
                n_inliers[f] = 0; iters_run[f] = 0;
            }
        }
        return;
    }

    // ---- 1. winner ---------------------------------------------------------------------------
    if (warp == 0) {
        if (n == 5) {
            if (tid == 0) {
                // cv2 skips RANSAC when npoints == model_points: direct solve, every point is an inlier
                int idx[5] = {0, 1, 2, 3, 4};
                double rv[3], tv[3];
                solve_minimal(X, uv, idx, cam, rv, tv);
                for (int k = 0; k < 3; ++k) { S.hyp_rt[k] = rv[k]; S.hyp_rt[3 + k] = tv[k]; }
                S.best_hyp = -1; S.iters = 0; S.have_best = 0;
            }
        } else {
            // the sequential loop `for (it = 0; it < niters && it < H; ++it)` of cv2 replayed by warp 0 (replay_accept); it leaves
            // the loop at max(niters, last accepted + 1), capped by H
            int niters = max_iters, best_count = 0, best_pos = -1;
            replay_accept(hyp_counts + (long long)f * H, 0, H, n, confidence, niters, best_count, best_pos);
            if (tid == 0) {
                int it = niters > best_pos + 1 ? niters : best_pos + 1;
                S.iters = it < H ? it : H;
                S.best_hyp = best_pos; S.have_best = 0;
                if (best_pos >= 0) for (int k = 0; k < 6; ++k) S.hyp_rt[k] = hyp_pose[((long long)f * H + best_pos) * 6 + k];
            }
        }
    }
    __syncthreads();
    if (n == 5) {
        if (tid == 0) {
            double R[9];
            rodrigues_to_mat(S.hyp_rt, R);
            for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) out[r * 4 + c] = R[3 * r + c]; out[r * 4 + 3] = S.hyp_rt[3 + r] / scale; }
            n_inliers[f] = 5; iters_run[f] = 0;
        }
        if (tid < 5) mask[tid] = 1;
        return;
    }
    if (S.best_hyp < 0) { write_identity(); if (tid == 0) iters_run[f] = S.iters; return; }

    // ---- 2. inlier mask of the winner + ordered compaction --------------------------------------
    {
        double R[9];
        rodrigues_to_mat(S.hyp_rt, R);
        if (tid == 0) S.base = 0;
        __syncthreads();
        for (int start = 0; start < n; start += blockDim.x) {
            const int i = start + tid;
            const bool in = i < n && reproj_err(X + 3 * i, uv + 2 * i, R, S.hyp_rt + 3, cam) <= thr2;
            const unsigned bal = __ballot_sync(0xffffffffu, in);
            if (lane == 0) S.warp_tot[warp] = __popc(bal);
            __syncthreads();
            int woff = 0, total = 0;
            for (int w = 0; w < kRefitThreads / 32; ++w) { if (w < warp) woff += S.warp_tot[w]; total += S.warp_tot[w]; }
            if (in) { mask[i] = 1; inl[S.base + woff + __popc(bal & ((1u << lane) - 1))] = i; }
            __syncthreads();
            if (tid == 0) S.base += total;
            __syncthreads();
        }
        if (tid == 0) S.n_in = S.base;
        __syncthreads();
    }
    const int m = S.n_in;
    const double ifx = 1. / cam.fx, ify = 1. / cam.fy;
    auto us_of = [&](int i, double &u, double &v) {      // float64 normalised coords (undistortPoints on CV_64F input)
        u = ((double)uv[2 * i] - cam.cx) * ifx; v = ((double)uv[2 * i + 1] - cam.cy) * ify;
    };

    // ---- 3. EPnP on all inliers: control points ---------------------------------------------------
    double acc[78];
    for (int k = 0; k < 3; ++k) acc[k] = 0;
    for (int q = tid; q < m; q += blockDim.x) { const float *p = X + 3 * inl[q]; acc[0] += p[0]; acc[1] += p[1]; acc[2] += p[2]; }
    block_reduce<3>(acc, S.scratch, S.red);
    if (tid == 0) for (int k = 0; k < 3; ++k) { S.cws[0][k] = S.red[k] / m; S.pw0[k] = S.red[k] / m; }
    __syncthreads();
    for (int k = 0; k < 6; ++k) acc[k] = 0;
    for (int q = tid; q < m; q += blockDim.x) {
        const float *p = X + 3 * inl[q];
        const double d0 = p[0] - S.cws[0][0], d1 = p[1] - S.cws[0][1], d2 = p[2] - S.cws[0][2];
        acc[0] += d0 * d0; acc[1] += d0 * d1; acc[2] += d0 * d2; acc[3] += d1 * d1; acc[4] += d1 * d2; acc[5] += d2 * d2;
    }
    block_reduce<6>(acc, S.scratch, S.red);
    if (tid == 0) {
        double c[9] = {S.red[0], S.red[1], S.red[2], S.red[1], S.red[3], S.red[4], S.red[2], S.red[4], S.red[5]};
        control_points_from_scatter(c, m, S.cws);
        double cc[9];
        for (int i = 0; i < 3; ++i) for (int j = 1; j < 4; ++j) cc[3 * i + j - 1] = S.cws[j][i] - S.cws[0][i];
        inv3(cc, S.ci);
    }
    __syncthreads();

    // ---- 4. barycentric coordinates + M^T M (78 upper-triangular sums) -------------------------------
    for (int k = 0; k < 78; ++k) acc[k] = 0;
    for (int q = tid; q < m; q += blockDim.x) {
        const int i = inl[q];
        const float *p = X + 3 * i;
        double a[4];
        for (int j = 0; j < 3; ++j)
            a[1 + j] = S.ci[3 * j] * (p[0] - S.cws[0][0]) + S.ci[3 * j + 1] * (p[1] - S.cws[0][1]) + S.ci[3 * j + 2] * (p[2] - S.cws[0][2]);
        a[0] = 1.0 - a[1] - a[2] - a[3];
        for (int j = 0; j < 4; ++j) alphas[4 * q + j] = a[j];
        double u, v;
        us_of(i, u, v);
        double m1[12], m2[12];
        for (int k = 0; k < 4; ++k) {
            m1[3 * k] = a[k]; m1[3 * k + 1] = 0.0; m1[3 * k + 2] = a[k] * (0.0 - u);
            m2[3 * k] = 0.0; m2[3 * k + 1] = a[k]; m2[3 * k + 2] = a[k] * (0.0 - v);
        }
        int e = 0;
#pragma unroll
        for (int r = 0; r < 12; ++r)
#pragma unroll
            for (int s = r; s < 12; ++s) acc[e++] += m1[r] * m1[s] + m2[r] * m2[s];
    }
    block_reduce<78>(acc, S.scratch, S.red);
    if (tid == 0) {
        int e = 0;
        for (int r = 0; r < 12; ++r) for (int s = r; s < 12; ++s) { S.ut[12 * r + s] = S.red[e]; S.ut[12 * s + r] = S.red[e]; ++e; }
        double dsv[12];
        jacobi_svd<12, 12, false>(S.ut, dsv, nullptr);
        solve_betas(S.ut, S.cws, S.betas);
    }
    __syncthreads();

    // ---- 5. three candidates: camera-frame points, Procrustes, reprojection error --------------------
    for (int cand = 0; cand < 3; ++cand) {
        if (tid == 0) compute_ccs(S.ut, S.betas[cand], S.ccs);
        __syncthreads();
        const double *a0 = alphas;     // sign test uses the first correspondence (solve_for_sign)
        const double z0 = a0[0] * S.ccs[0][2] + a0[1] * S.ccs[1][2] + a0[2] * S.ccs[2][2] + a0[3] * S.ccs[3][2];
        const double sgn = z0 < 0.0 ? -1.0 : 1.0;
        auto pc_of = [&](int q, double *pc) {
            const double *a = alphas + 4 * q;
            for (int j = 0; j < 3; ++j) pc[j] = sgn * (a[0] * S.ccs[0][j] + a[1] * S.ccs[1][j] + a[2] * S.ccs[2][j] + a[3] * S.ccs[3][j]);
        };
        for (int k = 0; k < 3; ++k) acc[k] = 0;
        for (int q = tid; q < m; q += blockDim.x) { double pc[3]; pc_of(q, pc); acc[0] += pc[0]; acc[1] += pc[1]; acc[2] += pc[2]; }
        block_reduce<3>(acc, S.scratch, S.red);
        const double pc0[3] = {S.red[0] / m, S.red[1] / m, S.red[2] / m};
        __syncthreads();
        for (int k = 0; k < 9; ++k) acc[k] = 0;
        for (int q = tid; q < m; q += blockDim.x) {
            double pc[3]; pc_of(q, pc);
            const float *pw = X + 3 * inl[q];
            for (int j = 0; j < 3; ++j) {
                acc[3 * j] += (pc[j] - pc0[j]) * (pw[0] - S.pw0[0]);
                acc[3 * j + 1] += (pc[j] - pc0[j]) * (pw[1] - S.pw0[1]);
                acc[3 * j + 2] += (pc[j] - pc0[j]) * (pw[2] - S.pw0[2]);
            }
        }
        block_reduce<9>(acc, S.scratch, S.red);
        if (tid == 0) rotation_from_abt(S.red, pc0, S.pw0, S.R, S.t);
        __syncthreads();
        acc[0] = 0;
        for (int q = tid; q < m; q += blockDim.x) {
            const int i = inl[q];
            const double pw[3] = {X[3 * i], X[3 * i + 1], X[3 * i + 2]};
            const double Xc = dot3(S.R, pw) + S.t[0], Yc = dot3(S.R + 3, pw) + S.t[1], inv_Zc = 1.0 / (dot3(S.R + 6, pw) + S.t[2]);
            const double ue = Xc * inv_Zc, ve = Yc * inv_Zc;
            double u, v;
            us_of(i, u, v);
            acc[0] += sqrt((u - ue) * (u - ue) + (v - ve) * (v - ve));
        }
        block_reduce<1>(acc, S.scratch, S.red);
        if (tid == 0) {
            const double rep = S.red[0] / m;
            if (!S.have_best || rep < S.best_rep) {
                S.have_best = 1; S.best_rep = rep;
                for (int k = 0; k < 9; ++k) S.best_R[k] = S.R[k];
                for (int k = 0; k < 3; ++k) S.best_t[k] = S.t[k];
            }
        }
        __syncthreads();
    }
    if (tid == 0) {
        double rvec[3], R[9];
        rodrigues_to_vec(S.best_R, rvec);
        rodrigues_to_mat(rvec, R);                    // the wrapper's cv2.Rodrigues(rvec), eval_utils.py:31
        for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) out[r * 4 + c] = R[3 * r + c]; out[r * 4 + 3] = S.best_t[r] / scale; }
        n_inliers[f] = m; iters_run[f] = S.iters;
    }
}

constexpr int kRoundCap = 4096;          // hypotheses per adaptive round at most (sizes round_rt)
struct PnpWorkspace { float *X32, *uv32; double *alphas, *round_rt; int32_t *inl, *bound, *state, *samples, *sample_set; unsigned long long *rng; size_t bytes; };
static PnpWorkspace carve_pnp(void *base, int F, int stride, int H) {
    PnpWorkspace w; size_t off = 0;
    auto take = [&](size_t bytes) { void *p = base ? (char *)base + off : nullptr; off += align_up(bytes, 256); return p; };
    w.alphas = (double *)take((size_t)F * stride * 4 * sizeof(double));
    w.X32 = (float *)take((size_t)F * stride * 3 * sizeof(float));
    w.uv32 = (float *)take((size_t)F * stride * 2 * sizeof(float));
    w.inl = (int32_t *)take((size_t)F * stride * sizeof(int32_t));
    w.bound = (int32_t *)take((size_t)F * sizeof(int32_t));
    w.state = (int32_t *)take((size_t)F * 2 * sizeof(int32_t));
    w.rng = (unsigned long long *)take((size_t)F * sizeof(unsigned long long));
    w.sample_set = (int32_t *)take((size_t)F * sizeof(int32_t));
    w.samples = (int32_t *)take((size_t)F * H * 5 * sizeof(int32_t));     // cv2's stream when the caller supplies no samples
    w.round_rt = (double *)take((size_t)F * (H < kRoundCap ? H : kRoundCap) * 12 * sizeof(double));   // [R | t] of one round's hypotheses
    w.bytes = off;
    return w;
}

}  // namespace pnp
}  // namespace opb

extern "C" size_t op_pnp_workspace_bytes(int F, int stride, int H) {
    return opb::pnp::carve_pnp(nullptr, F > 0 ? F : 1, stride > 0 ? stride : 1, H > 0 ? H : 1).bytes;
}

extern "C" int op_pnp_cv_samples(const int32_t *count, int F, int H, int32_t *samples, int32_t *sample_set, void *stream) {
    OPB_CHECK_ARG(count && samples && sample_set && F > 0 && H > 0, "op_pnp_cv_samples: bad argument");
    opb::pnp::k_pnp_cv_samples<<<opb::ceil_div(F, 32), 32, 0, (cudaStream_t)stream>>>(count, F, H, 0, H, nullptr, nullptr, samples, sample_set);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_pnp_ransac(const double *pts2d, const double *pts3d, const int32_t *count, int F, int stride,
                             const double *K, const int32_t *samples, const int32_t *sample_set, int H,
                             double reproj_thr, double confidence, int max_iters, double scale,
                             double *pose, uint8_t *inlier_mask, int32_t *n_inliers, int32_t *iters_run,
                             int32_t *hyp_counts, double *hyp_pose, void *workspace, size_t workspace_bytes, void *stream) {
    using namespace opb;
    using namespace opb::pnp;
    OPB_CHECK_ARG(pts2d && pts3d && count && K && pose && inlier_mask && n_inliers && iters_run &&
                  hyp_counts && hyp_pose && workspace, "op_pnp_ransac: null pointer");
    OPB_CHECK_ARG((samples == nullptr) == (sample_set == nullptr), "op_pnp_ransac: samples and sample_set come together (both NULL = cv2's own stream)");
    OPB_CHECK_ARG(F > 0 && stride > 0 && H > 0 && F <= 65535, "op_pnp_ransac: bad shape F=%d stride=%d H=%d", F, stride, H);
    OPB_CHECK_ARG(scale != 0.0 && reproj_thr > 0, "op_pnp_ransac: bad scale / threshold");
    PnpWorkspace w = carve_pnp(workspace, F, stride, H);
    if (workspace_bytes < w.bytes) { set_error("op_pnp_ransac: workspace %zu < %zu bytes", workspace_bytes, w.bytes); return OPB_ERR_WORKSPACE; }
    cudaStream_t st = (cudaStream_t)stream;
    const float thr2 = (float)(reproj_thr * reproj_thr);
    k_pnp_prepare<<<dim3(ceil_div(stride, 256), F), 256, 0, st>>>(pts2d, pts3d, count, stride, w.X32, w.uv32);
    OPB_CHECK_LAUNCH();
    // Adaptive rounds (exactly cv2's result): round r evaluates hypotheses [h0, h1) only for frames whose iteration bound --
    // cv2's niters replayed over the rounds before -- still reaches h0.  Unevaluated hyp_counts stay -1.
    OPB_CHECK_CUDA(cudaMemsetAsync(hyp_counts, 0xFF, (size_t)F * H * sizeof(int32_t), st));
    const bool own_stream = samples == nullptr;
    if (own_stream) { samples = w.samples; sample_set = w.sample_set; }
    static const int kRoundEnd[] = {32, 256, 1024, 4096};
    auto round_end = [&](int r, int h0) { int h1 = r < 4 ? kRoundEnd[r] : h0 + 4096; return h1 > H ? H : h1; };
    const int cap = H < kRoundCap ? H : kRoundCap;
    int h0 = 0;
    for (int r = 0; h0 < H; ++r) {
        const int h1 = round_end(r, h0);
        const int32_t *bnd = r == 0 ? nullptr : w.bound;
        NextSamples nx{};
        if (own_stream) {
            if (r == 0) {
                k_pnp_cv_samples<<<ceil_div(F, 32), 32, 0, st>>>(count, F, H, h0, h1, bnd, w.rng, w.samples, w.sample_set);
                OPB_CHECK_LAUNCH();
            }
            // the next round's samples are drawn by one extra block column of this round's solve
            if (h1 < H) nx = NextSamples{F, h1, round_end(r + 1, h1), w.rng, w.samples, w.sample_set};
        }
        // two blocks per SM at 255 registers (default); OPB_PNP_SOLVE_OCC=3 / 4 trade registers (168 / 128, more spills) for resident warps:
        // measured slower at configs[3] (8.6 vs 9.1 / 9.2 ms per 64 frames), kept for A/B runs
        static const int occ = [] { const char *e = getenv("OPB_PNP_SOLVE_OCC"); return e ? atoi(e) : 2; }();
        const dim3 sgrid(ceil_div(h1 - h0, kHypBlock) + (nx.h1 > nx.h0 ? 1 : 0), F);
        if (occ == 4)
            k_pnp_solve<4><<<sgrid, kHypBlock, 0, st>>>(w.X32, w.uv32, count, stride, K, samples, sample_set, H, h0, h1, cap, bnd, hyp_counts, hyp_pose, w.round_rt, nx);
        else if (occ == 3)
            k_pnp_solve<3><<<sgrid, kHypBlock, 0, st>>>(w.X32, w.uv32, count, stride, K, samples, sample_set, H, h0, h1, cap, bnd, hyp_counts, hyp_pose, w.round_rt, nx);
        else
            k_pnp_solve<2><<<sgrid, kHypBlock, 0, st>>>(w.X32, w.uv32, count, stride, K, samples, sample_set, H, h0, h1, cap, bnd, hyp_counts, hyp_pose, w.round_rt, nx);
        k_pnp_count<<<dim3(ceil_div(h1 - h0, kCountWarps), F), kCountWarps * 32, 0, st>>>(w.X32, w.uv32, count, stride, K, H, h0, h1, cap, bnd,
                                                                                         thr2, w.round_rt, hyp_counts);
        OPB_CHECK_LAUNCH();
        if (h1 < H) {
            k_pnp_bound_update<<<F, 32, 0, st>>>(count, F, H, h0, h1, hyp_counts, confidence, max_iters, w.bound, w.state);
            OPB_CHECK_LAUNCH();
        }
        h0 = h1;
    }
    k_pnp_select_refit<<<F, kRefitThreads, 0, st>>>(w.X32, w.uv32, count, stride, K, hyp_counts, hyp_pose, H, thr2, confidence,
                                                    max_iters, scale, w.alphas, w.inl, pose, inlier_mask, n_inliers, iters_run);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
