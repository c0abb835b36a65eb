// (4) Batched PnP-RANSAC -- replaces cv2.solvePnPRansac(..., reprojectionError=5, iterationsCount=10000,
// flags=SOLVEPNP_EPNP) as called by ransac_PnP (src/utils/eval_utils.py:18-42).
//
//  k_pnp_prepare     : float64 correspondences -> float32 (cv2 converts its inputs to CV_32F first).
//  k_pnp_solve       : lane-per-hypothesis 5-point EPnP (a warp solves 32 minimal samples at once, every lane busy).
//  k_pnp_count       : warp-per-hypothesis inlier count (lanes over correspondences): projection in float64,
//                      squared error in float32, inlier iff err <= thr^2 -- the arithmetic of
//                      PnPRansacCallback::computeError.
//  k_pnp_select_refit: replays cv2's sequential rule over the counts (accept iff strictly greater than
//                      every earlier count and > 4; niters shrinks adaptively; stop at iter == niters),
//                      recomputes the winner's inlier mask, refits EPnP on all inliers with block-parallel
//                      ordered reductions, and writes [Rodrigues(rvec) | t / scale].
// Whole file is compiled with -fmad=false (see onepose_b200/build.py).
#include "common.cuh"
#include "pnp_epnp.cuh"

namespace opb {
namespace pnp {

constexpr int kHypBlock = 128;
constexpr int kRefitThreads = 256;

struct Cam { double fx, fy, cx, cy; };
__device__ __forceinline__ Cam load_cam(const double *K) { return Cam{K[0], K[4], K[2], K[5]}; }

__global__ void k_pnp_prepare(const double *__restrict__ pts2d, const double *__restrict__ pts3d, const int32_t *__restrict__ count,
                              int stride, float *__restrict__ X32, float *__restrict__ uv32) {
    const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count[f]) return;
    const long long o = (long long)f * stride + i;
    X32[o * 3] = (float)pts3d[o * 3]; X32[o * 3 + 1] = (float)pts3d[o * 3 + 1]; X32[o * 3 + 2] = (float)pts3d[o * 3 + 2];
    uv32[o * 2] = (float)pts2d[o * 2]; uv32[o * 2 + 1] = (float)pts2d[o * 2 + 1];
}

// squared reprojection error in float32, exactly as cv2: project in float64, round the pixel to float32,
// difference / square / add in float32
__device__ __forceinline__ float reproj_err(const float *X, const float *uv, const double *R, const double *t, const Cam &cam) {
    const double Xd = X[0], Yd = X[1], Zd = X[2];
    double x = R[0] * Xd + R[1] * Yd + R[2] * Zd + t[0];
    double y = R[3] * Xd + R[4] * Yd + R[5] * Zd + t[1];
    double z = R[6] * Xd + R[7] * Yd + R[8] * Zd + t[2];
    z = z != 0.0 ? 1. / z : 1.0;
    x *= z; y *= z;
    const float pu = (float)(x * cam.fx + cam.cx), pv = (float)(y * cam.fy + cam.cy);
    const float d0 = uv[0] - pu, d1 = uv[1] - pv;
    float s = d0 * d0;
    s += d1 * d1;
    return s;
}

// 5-point hypothesis from float32 points (PnPRansacCallback::runKernel -> solvePnP(EPNP)):
// normalised image coordinates are rounded to float32 by undistortPoints, EPnP runs with the identity camera.
__device__ inline void solve_minimal(const float *X32, const float *uv32, const int *idx, const Cam &cam, double *rvec, double *tvec) {
    const double ifx = 1. / cam.fx, ify = 1. / cam.fy;
    double pws[15], us[10], R[9];
    for (int i = 0; i < 5; ++i) {
        for (int k = 0; k < 3; ++k) pws[3 * i + k] = X32[3 * idx[i] + k];
        const float xn = (float)(((double)uv32[2 * idx[i]] - cam.cx) * ifx);
        const float yn = (float)(((double)uv32[2 * idx[i] + 1] - cam.cy) * ify);
        us[2 * i] = xn; us[2 * i + 1] = yn;
    }
    epnp_small<5>(pws, us, R, tvec);
    rodrigues_to_vec(R, rvec);
}

__device__ inline void cv_samples_frame(int f, const int32_t *__restrict__ count, int H, int h_begin, int h_end,
                                        const int32_t *__restrict__ bound, unsigned long long *__restrict__ rng_state,
                                        int32_t *__restrict__ samples, int32_t *__restrict__ sample_set);

// The sequential sample stream of the NEXT round rides along as one extra block column (blockIdx.x == gridDim.x - 1 when next_h1 >
// next_h0): it only depends on the frames' counts, and on its own launch it was 0.07-0.9 ms per round of one thread per frame with
// the rest of the GPU idle.
struct NextSamples { int F, h0, h1; unsigned long long *rng; int32_t *samples, *sample_set; };

// Hypothesis round, stage 1: lane-per-hypothesis minimal solve.  Writes (rvec, tvec) to hyp_pose, the rotation matrix cv2's
// projectPoints re-derives from rvec plus tvec to round_rt [F][round_cap][12] (slot h - h_begin), and marks the hypothesis in
// hyp_counts: 0 = solved, to be counted by k_pnp_count; -1 = not evaluated (bad sample, n <= 5, or past the adaptive bound).
template <int MIN_BLOCKS>
__global__ void __launch_bounds__(kHypBlock, MIN_BLOCKS)
k_pnp_solve(const float *__restrict__ X32, const float *__restrict__ uv32, const int32_t *__restrict__ count, int stride,
            const double *__restrict__ Kmat, const int32_t *__restrict__ samples, const int32_t *__restrict__ sample_set,
            int H_total, int h_begin, int h_end, int round_cap, const int32_t *__restrict__ bound, int32_t *__restrict__ hyp_counts,
            double *__restrict__ hyp_pose, double *__restrict__ round_rt, const NextSamples nx) {
    if (nx.h1 > nx.h0 && blockIdx.x == gridDim.x - 1) {
        if (blockIdx.y == 0)
            for (int g = threadIdx.x; g < nx.F; g += kHypBlock) cv_samples_frame(g, count, H_total, nx.h0, nx.h1, bound, nx.rng, nx.samples, nx.sample_set);
        return;
    }
    const int f = blockIdx.y, lane = threadIdx.x & 31;
    const int h = h_begin + blockIdx.x * kHypBlock + threadIdx.x;
    // adaptive rounds: cv2's sequential loop never looks at iteration >= niters, and niters only shrinks, so a warp whose
    // first hypothesis is already past the bound replayed over the earlier rounds has nothing to contribute
    if (h - lane >= h_end || (bound && h - lane >= bound[f])) return;
    const int H = h_end < H_total ? h_end : H_total;
    if (h >= H) return;
    const int n = count[f];
    if (n <= 5) { hyp_counts[(long long)f * H_total + h] = -1; return; }       // no RANSAC for n <= 5 (handled in k_pnp_select_refit)
    const Cam cam = load_cam(Kmat + (long long)f * 9);
    const float *X = X32 + (long long)f * stride * 3, *uv = uv32 + (long long)f * stride * 2;
    const int32_t *s = samples + ((long long)sample_set[f] * H_total + h) * 5;
    int idx[5];
    bool ok = true;
    for (int i = 0; i < 5; ++i) { idx[i] = s[i]; ok = ok && idx[i] >= 0 && idx[i] < n; }
    if (!ok) { hyp_counts[(long long)f * H_total + h] = -1; return; }
    double rvec[3], tvec[3], R[9];
    solve_minimal(X, uv, idx, cam, rvec, tvec);
    double *dst = hyp_pose + ((long long)f * H_total + h) * 6;
    dst[0] = rvec[0]; dst[1] = rvec[1]; dst[2] = rvec[2]; dst[3] = tvec[0]; dst[4] = tvec[1]; dst[5] = tvec[2];
    rodrigues_to_mat(rvec, R);          // cv2 re-derives R from rvec inside projectPoints
    double *rt = round_rt + ((long long)f * round_cap + (h - h_begin)) * 12;
#pragma unroll
    for (int k = 0; k < 9; ++k) rt[k] = R[k];
    rt[9] = tvec[0]; rt[10] = tvec[1]; rt[11] = tvec[2];
    hyp_counts[(long long)f * H_total + h] = 0;
}

// Hypothesis round, stage 2: inlier count, one WARP per hypothesis (lanes over the frame's correspondences) -- frames x hypotheses
// warps fill the machine where the former one-warp-walks-its-32-poses form left 148 SMs to a few warps per frame (BASELINE.json
// configs[3]: 10k correspondences).  Projection in float64, squared error in float32, inlier iff err <= thr^2: the arithmetic of
// PnPRansacCallback::computeError; the count is an integer sum, so the split over lanes does not change it.
constexpr int kCountWarps = 8, kCountChunk = 1024;
// the same arithmetic on a point already widened to float64 (the widening is exact)
__device__ __forceinline__ float reproj_err_d(const double *Xd, const float2 uv, const double *R, const double *t, const Cam &cam) {
    double x = R[0] * Xd[0] + R[1] * Xd[1] + R[2] * Xd[2] + t[0];
    double y = R[3] * Xd[0] + R[4] * Xd[1] + R[5] * Xd[2] + t[1];
    double z = R[6] * Xd[0] + R[7] * Xd[1] + R[8] * Xd[2] + t[2];
    z = z != 0.0 ? 1. / z : 1.0;
    x *= z; y *= z;
    const float pu = (float)(x * cam.fx + cam.cx), pv = (float)(y * cam.fy + cam.cy);
    const float d0 = uv.x - pu, d1 = uv.y - pv;
    float s = d0 * d0;
    s += d1 * d1;
    return s;
}
// A block = 8 consecutive hypotheses of one frame.  The frame's correspondences pass through shared memory in chunks of 1024,
// widened to float64 once per block instead of once per hypothesis (the first form read them from global memory inside the loop:
// ncu long_scoreboard 0.8 of the samples); double[3] rows (stride 24 B) and float2 pixels are bank-conflict free.
__global__ void __launch_bounds__(kCountWarps * 32)
k_pnp_count(const float *__restrict__ X32, const float *__restrict__ uv32, const int32_t *__restrict__ count, int stride,
            const double *__restrict__ Kmat, int H_total, int h_begin, int h_end, int round_cap, const int32_t *__restrict__ bound,
            float thr2, const double *__restrict__ round_rt, int32_t *__restrict__ hyp_counts) {
    __shared__ double sX[kCountChunk * 3];
    __shared__ float2 suv[kCountChunk];
    const int f = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int hb = h_begin + blockIdx.x * kCountWarps, h = hb + warp;
    const int H = h_end < H_total ? h_end : H_total;
    if (hb >= H || (bound && (hb & ~31) >= bound[f])) return;               // block-uniform: the bound test of the solving warp (32 consecutive hypotheses)
    const bool active = h < H && hyp_counts[(long long)f * H_total + h] >= 0;   // solved by k_pnp_solve
    const int n = count[f];
    const Cam cam = load_cam(Kmat + (long long)f * 9);
    const float *X = X32 + (long long)f * stride * 3;
    const float2 *uv = reinterpret_cast<const float2 *>(uv32 + (long long)f * stride * 2);
    double R[9], t[3];
    if (active) {
        const double *rt = round_rt + ((long long)f * round_cap + (h - h_begin)) * 12;
#pragma unroll
        for (int k = 0; k < 9; ++k) R[k] = rt[k];
        t[0] = rt[9]; t[1] = rt[10]; t[2] = rt[11];
    }
    int c = 0;
    for (int c0 = 0; c0 < n; c0 += kCountChunk) {
        const int m = min(kCountChunk, n - c0);
        for (int i = threadIdx.x; i < 3 * m; i += kCountWarps * 32) sX[i] = (double)X[3 * c0 + i];
        for (int i = threadIdx.x; i < m; i += kCountWarps * 32) suv[i] = uv[c0 + i];
        __syncthreads();
        if (active)
            for (int i = lane; i < m; i += 32) c += reproj_err_d(sX + 3 * i, suv[i], R, t, cam) <= thr2 ? 1 : 0;
        __syncthreads();
    }
    if (!active) return;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) hyp_counts[(long long)f * H_total + h] = c;
}

// cv::RNG((uint64)-1) multiply-with-carry stream, one sequential generator per frame.  Samples [h_begin, h_end) are
// drawn; the generator state is carried in rng_state between rounds (rng_state == nullptr: one shot from the seed).
__device__ inline void cv_samples_frame(int f, const int32_t *__restrict__ count, int H, int h_begin, int h_end,
                                        const int32_t *__restrict__ bound, unsigned long long *__restrict__ rng_state,
                                        int32_t *__restrict__ samples, int32_t *__restrict__ sample_set) {
    if (h_begin == 0) sample_set[f] = f;
    if (bound && h_begin >= bound[f]) return;
    const unsigned n = (unsigned)count[f];
    int32_t *out = samples + (long long)f * H * 5;
    if (n < 5) { for (int i = h_begin * 5; i < h_end * 5; ++i) out[i] = 0; return; }
    unsigned long long state = (h_begin == 0 || !rng_state) ? 0xffffffffffffffffull : rng_state[f];
    for (int it = h_begin; it < h_end; ++it) {
        int idx[5];
        for (int i = 0; i < 5; ++i) {
            int v;
            for (;;) {
                state = (unsigned long long)(unsigned)state * 4164903690ull + (state >> 32);
                v = (int)((unsigned)state % n);
                bool dup = false;
                for (int k = 0; k < i; ++k) dup = dup || idx[k] == v;
                if (!dup) break;
            }
            idx[i] = v;
            out[it * 5 + i] = v;
        }
    }
    if (rng_state) rng_state[f] = state;
}

__global__ void k_pnp_cv_samples(const int32_t *__restrict__ count, int F, int H, int h_begin, int h_end,
                                 const int32_t *__restrict__ bound, unsigned long long *__restrict__ rng_state,
                                 int32_t *__restrict__ samples, int32_t *__restrict__ sample_set) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < F) cv_samples_frame(f, count, H, h_begin, h_end, bound, rng_state, samples, sample_set);
}

__device__ inline int update_num_iters(double p, double ep, int model_points, int max_iters) {
    p = p < 0 ? 0 : p > 1 ? 1 : p;
    ep = ep < 0 ? 0 : ep > 1 ? 1 : ep;
    double num = 1 - p > DBL_MIN ? 1 - p : DBL_MIN;
    double denom = 1 - pow(1 - ep, (double)model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num); denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)llrint(num / denom);
}

// cv2's sequential accept rule over the counts of iterations [it_begin, it_end), run by one warp (all 32 lanes call with the same
// arguments and leave with the same results).  The rule -- accept iff count > max(best so far, 4); niters shrinks on every accept;
// the loop ends at it == niters -- only changes state at RECORD positions (count above everything before it), so the warp reads the
// counts 32 at a time, finds the records with a prefix maximum and walks only those in order; a one-thread scan of the list was a
// chain of dependent global loads (0.25 ms for 3072 iterations).  Unevaluated entries (-1) are never records.
__device__ inline void replay_accept(const int32_t *__restrict__ cnt, int it_begin, int it_end, int n, double confidence,
                                     int &niters, int &best, int &best_pos) {
    const int lane = threadIdx.x & 31;
    for (int it0 = it_begin; it0 < it_end && it0 < niters; it0 += 32) {
        const int it = it0 + lane;
        const int good = (it < it_end) ? cnt[it] : -1;
        int pm = good;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, pm, o); if (lane >= o) pm = pm > t ? pm : t; }
        int excl = __shfl_up_sync(0xffffffffu, pm, 1);
        if (lane == 0) excl = -1;
        const int floor_ = best > 4 ? best : 4;
        unsigned rec = __ballot_sync(0xffffffffu, good > (excl > floor_ ? excl : floor_));
        while (rec) {
            const int l = __ffs(rec) - 1;
            rec &= rec - 1;
            const int pos = it0 + l;
            if (pos >= niters) { rec = 0; break; }                     // the sequential loop ended before this iteration
            best = __shfl_sync(0xffffffffu, good, l); best_pos = pos;
            niters = update_num_iters(confidence, (double)(n - best) / n, 5, niters);
        }
    }
}

// Tightens the per-frame iteration bound after round [h_begin, h_end) (state = {niters, best_count}).  One warp per frame.
__global__ void __launch_bounds__(32)
k_pnp_bound_update(const int32_t *__restrict__ count, int F, int H, int h_begin, int h_end,
                   const int32_t *__restrict__ hyp_counts, double confidence, int max_iters,
                   int32_t *__restrict__ bound, int32_t *__restrict__ state) {
    const int f = blockIdx.x;
    int niters = h_begin == 0 ? max_iters : state[2 * f], best = h_begin == 0 ? 0 : state[2 * f + 1], pos = -1;
    const int n = count[f];
    if (n > 5) replay_accept(hyp_counts + (long long)f * H, h_begin, h_end, n, confidence, niters, best, pos);
    else niters = 0;
    if (threadIdx.x == 0) { state[2 * f] = niters; state[2 * f + 1] = best; bound[f] = niters; }
}

// ordered block reduction of K doubles per thread; result broadcast to all threads through `out`
template <int K>
__device__ void block_reduce(double *v, double *scratch /* [8][K] */, double *out /* [K] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) scratch[warp * K + k] = x;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        double s = 0;
        for (int w = 0; w < kRefitThreads / 32; ++w) s += scratch[w * K + k];
        out[k] = s;
    }
    __syncthreads();
}

}  // namespace pnp
}  // namespace opb

#include "pnp_p3p.cuh"
#include "pnp_refit.cuh"
#include "pnp_lm.cuh"
