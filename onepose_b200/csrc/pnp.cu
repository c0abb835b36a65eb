// (4) Batched PnP-RANSAC -- replaces cv2.solvePnPRansac(..., reprojectionError=5, iterationsCount=10000,
// flags=SOLVEPNP_EPNP) as called by ransac_PnP (src/utils/eval_utils.py:18-42).
//
//  k_pnp_prepare     : float64 correspondences -> float32 (cv2 converts its inputs to CV_32F first).
//  k_pnp_hypotheses  : lane-per-hypothesis 5-point EPnP (a warp solves 32 minimal samples at once, every
//                      lane busy), then the same warp walks its 32 poses and counts inliers cooperatively
//                      (lanes over correspondences): projection in float64, squared error in float32,
//                      inlier iff err <= thr^2 -- the arithmetic of PnPRansacCallback::computeError.
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

__global__ void __launch_bounds__(kHypBlock)
k_pnp_hypotheses(const float *__restrict__ X32, const float *__restrict__ uv32, const int32_t *__restrict__ count, int stride,
                 const double *__restrict__ Kmat, const int32_t *__restrict__ samples, const int32_t *__restrict__ sample_set,
                 int H_total, int h_begin, int h_end, const int32_t *__restrict__ bound, float thr2, int32_t *__restrict__ hyp_counts,
                 double *__restrict__ hyp_pose) {
    const int f = blockIdx.y, lane = threadIdx.x & 31;
    const int h = h_begin + blockIdx.x * kHypBlock + threadIdx.x;
    // adaptive rounds: cv2's sequential loop never looks at iteration >= niters, and niters only shrinks, so a warp whose
    // first hypothesis is already past the bound replayed over the earlier rounds has nothing to contribute
    if (h - lane >= h_end || (bound && h - lane >= bound[f])) return;
    const int H = h_end < H_total ? h_end : H_total;
    const int n = count[f];
    if (n <= 5) {                       // no RANSAC for n <= 5 (handled in k_pnp_select_refit)
        if (h < H) hyp_counts[(long long)f * H_total + h] = -1;
        return;
    }
    const Cam cam = load_cam(Kmat + (long long)f * 9);
    const float *X = X32 + (long long)f * stride * 3, *uv = uv32 + (long long)f * stride * 2;
    double rvec[3] = {0, 0, 0}, tvec[3] = {0, 0, 0}, R[9];
    bool ok = false;
    if (h < H) {
        const int32_t *s = samples + ((long long)sample_set[f] * H_total + h) * 5;
        int idx[5];
        ok = true;
        for (int i = 0; i < 5; ++i) { idx[i] = s[i]; ok = ok && idx[i] >= 0 && idx[i] < n; }
        if (ok) {
            solve_minimal(X, uv, idx, cam, rvec, tvec);
            double *dst = hyp_pose + ((long long)f * H_total + h) * 6;
            dst[0] = rvec[0]; dst[1] = rvec[1]; dst[2] = rvec[2]; dst[3] = tvec[0]; dst[4] = tvec[1]; dst[5] = tvec[2];
        }
    }
    rodrigues_to_mat(rvec, R);          // cv2 re-derives R from rvec inside projectPoints
    int my_count = -1;
    for (int hh = 0; hh < 32; ++hh) {
        const bool okh = __shfl_sync(0xffffffffu, (int)ok, hh) != 0;
        if (!okh) continue;             // warp-uniform
        double Rb[9], tb[3];
#pragma unroll
        for (int k = 0; k < 9; ++k) Rb[k] = __shfl_sync(0xffffffffu, R[k], hh);
#pragma unroll
        for (int k = 0; k < 3; ++k) tb[k] = __shfl_sync(0xffffffffu, tvec[k], hh);
        int c = 0;
        for (int i = lane; i < n; i += 32) c += reproj_err(X + 3 * i, uv + 2 * i, Rb, tb, cam) <= thr2 ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == hh) my_count = c;
    }
    if (h < H) hyp_counts[(long long)f * H_total + h] = my_count;
}

// cv::RNG((uint64)-1) multiply-with-carry stream, one sequential generator per frame.  Samples [h_begin, h_end) are
// drawn; the generator state is carried in rng_state between rounds (rng_state == nullptr: one shot from the seed).
__global__ void k_pnp_cv_samples(const int32_t *__restrict__ count, int F, int H, int h_begin, int h_end,
                                 const int32_t *__restrict__ bound, unsigned long long *__restrict__ rng_state,
                                 int32_t *__restrict__ samples, int32_t *__restrict__ sample_set) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
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

__device__ inline int update_num_iters(double p, double ep, int model_points, int max_iters) {
    p = p < 0 ? 0 : p > 1 ? 1 : p;
    ep = ep < 0 ? 0 : ep > 1 ? 1 : ep;
    double num = 1 - p > DBL_MIN ? 1 - p : DBL_MIN;
    double denom = 1 - pow(1 - ep, (double)model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num); denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)llrint(num / denom);
}

// Replays cv2's accept rule over the counts of round [h_begin, h_end) and tightens the per-frame iteration bound
// (state = {niters, best_count}).  One thread per frame: the rule is a dependent scan.
__global__ void k_pnp_bound_update(const int32_t *__restrict__ count, int F, int H, int h_begin, int h_end,
                                   const int32_t *__restrict__ hyp_counts, double confidence, int max_iters,
                                   int32_t *__restrict__ bound, int32_t *__restrict__ state) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    int niters = h_begin == 0 ? max_iters : state[2 * f], best = h_begin == 0 ? 0 : state[2 * f + 1];
    const int n = count[f];
    if (n > 5) {
        const int32_t *cnt = hyp_counts + (long long)f * H;
        for (int it = h_begin; it < h_end && it < niters; ++it) {
            const int good = cnt[it];
            const int floor_ = best > 4 ? best : 4;
            if (good > floor_) { best = good; niters = update_num_iters(confidence, (double)(n - good) / n, 5, niters); }
        }
    } else {
        niters = 0;
    }
    state[2 * f] = niters; state[2 * f + 1] = best;
    bound[f] = niters;
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
