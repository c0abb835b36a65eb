// (2) Fused linear layer on the 5th-gen tensor cores -- the 1x1 Conv1d projections of MultiHeadedAttention / MLP /
// final_proj (GATs_SuperGlue.py:108-120,135-147,194-199,241).  Kernel in linear_tc.cuh; this file holds the
// elementwise producers of the split-precision planes and the C ABI.
#include <cstring>
#include <stdlib.h>
#include "common.cuh"
#include "linear_tc.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace opb {

// fp32 -> (hi, lo) fp16 planes of 64*x
__global__ void k_split_planes64(const float *__restrict__ x, __half *__restrict__ hi, __half *__restrict__ lo, long long n) {
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const float4 v = *reinterpret_cast<const float4 *>(x + i);
    __half h[4], l[4];
    tc::split_f16(v.x * lintc::kScale, h[0], l[0]); tc::split_f16(v.y * lintc::kScale, h[1], l[1]);
    tc::split_f16(v.z * lintc::kScale, h[2], l[2]); tc::split_f16(v.w * lintc::kScale, h[3], l[3]);
    *reinterpret_cast<uint2 *>(hi + i) = *reinterpret_cast<uint2 *>(h);
    *reinterpret_cast<uint2 *>(lo + i) = *reinterpret_cast<uint2 *>(l);
}

// planes of relu((h - mean[b,c]) * rstd[b,c])  -- InstanceNorm1d + ReLU between the two MLP convolutions (:144-146)
__global__ void k_norm_relu_split(const float *__restrict__ h, const float *__restrict__ mean, const float *__restrict__ rstd,
                                  __half *__restrict__ hi, __half *__restrict__ lo, long long per_batch, int C, long long n) {
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const int b = (int)(i / per_batch), c = (int)(i % C);
    const float4 v = *reinterpret_cast<const float4 *>(h + i);
    const float4 mu = *reinterpret_cast<const float4 *>(mean + (long long)b * C + c);
    const float4 rs = *reinterpret_cast<const float4 *>(rstd + (long long)b * C + c);
    const float o[4] = {fmaxf((v.x - mu.x) * rs.x, 0.f), fmaxf((v.y - mu.y) * rs.y, 0.f), fmaxf((v.z - mu.z) * rs.z, 0.f),
                        fmaxf((v.w - mu.w) * rs.w, 0.f)};
    __half hh[4], ll[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) tc::split_f16(o[q] * lintc::kScale, hh[q], ll[q]);
    *reinterpret_cast<uint2 *>(hi + i) = *reinterpret_cast<uint2 *>(hh);
    *reinterpret_cast<uint2 *>(lo + i) = *reinterpret_cast<uint2 *>(ll);
}

// mean / rstd from the per-(row tile, warp) fp32 partials of the PRE-BIAS result, folded in a fixed order in float64
__global__ void k_instnorm_finalize_tc(const float2 *__restrict__ part, const float *__restrict__ bias, int parts, int C, int rows,
                                       float eps, float *__restrict__ mean, float *__restrict__ rstd) {
    const int b = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s = 0, q = 0;
    for (int t = 0; t < parts; ++t) {
        const float2 v = part[((long long)b * parts + t) * C + c];
        s += (double)v.x; q += (double)v.y;
    }
    const double mu = s / rows;
    double var = q / rows - mu * mu;     // biased variance, InstanceNorm1d default; a constant bias does not change it
    if (var < 0) var = 0;
    mean[(long long)b * C + c] = (float)(mu + (bias ? (double)bias[c] : 0.0));
    rstd[(long long)b * C + c] = (float)(1.0 / sqrt(var + (double)eps));
}

// (sum, 0) over groups of 32 rows, laid out like the COLSUM partials of op_linear_tc: partial[b][row/32][c]
__global__ void k_colsum_partial(const float *__restrict__ x, int rows, int C, int parts, float2 *__restrict__ partial) {
    const int part = blockIdx.x, b = blockIdx.y;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float s = 0.f;
        const int r0 = part * 32, r1 = min(rows, r0 + 32);
        for (int r = r0; r < r1; ++r) s += x[((long long)b * rows + r) * C + c];
        partial[((long long)b * parts + part) * C + c] = make_float2(s, 0.f);
    }
}

// Analytic InstanceNorm mean of MLP layer 0 with the merge projection folded in (:119,135-147):
//   mean_rows(h) = W0a mean_rows(x) + P'_b mean_rows(Q'') + const, so the centred pre-activation is obtained with the per-frame
// bias  -(W0a mean_x + P'_b mean_q)  (the constants cancel).  W0aT is [256 k][512 c], P' planes [B or 1][512 c][256 j] (x64).
// One cluster of 8 CTAs per frame: CTA r folds the column-sum partials of input channels [64r, 64r+64) of [x | Q''] (4 slices
// of the parts per channel, fixed order), the 512 means are exchanged through distributed shared memory, and CTA r then
// produces output columns [64r, 64r+64): the W0a part with k cut into 4 slices, the P' part with one warp per row.
__global__ void __cluster_dims__(1, 8, 1) __launch_bounds__(256)
k_mlp_center_bias(const float2 *__restrict__ xpart, int xparts, const float2 *__restrict__ qpart, int qparts,
                  const float *__restrict__ W0aT, const __half *__restrict__ Ph, const __half *__restrict__ Pl, int p_batched,
                  int rows, const int *__restrict__ row_counts, float *__restrict__ bias_out) {
    __shared__ float red[4][64];
    __shared__ float mean_loc[64];
    __shared__ __align__(16) float mean_in[512];
    __shared__ float pdot[64];
    cg::cluster_group cluster = cg::this_cluster();
    const int b = blockIdx.x, r = blockIdx.y, tid = threadIdx.x, j = tid & 63, sl = tid >> 6;
    if (row_counts) rows = max(1, min(rows, __ldg(row_counts + b)));      // ragged batch: the mean is over this frame's valid rows
    {
        const float2 *src = r < 4 ? xpart : qpart;
        const int parts = r < 4 ? xparts : qparts;
        const int kk = (r & 3) * 64 + j;
        double s = 0;
#pragma unroll 8
        for (int t = sl; t < parts; t += 4) s += (double)__ldg(&src[((long long)b * parts + t) * 256 + kk].x);
        red[sl][j] = (float)s;
    }
    __syncthreads();
    if (tid < 64) mean_loc[tid] = (float)(((double)red[0][tid] + (double)red[1][tid] + (double)red[2][tid] + (double)red[3][tid]) / rows);
    cluster.sync();
    for (int k = tid; k < 512; k += 256) mean_in[k] = cluster.map_shared_rank(mean_loc, k >> 6)[k & 63];
    cluster.sync();                                       // nobody leaves (or reuses red) while its means are still being read
    const int c = r * 64 + j;
    float acc = 0.f;
#pragma unroll 16
    for (int k = sl * 64; k < sl * 64 + 64; ++k) acc = fmaf(__ldg(W0aT + (long long)k * 512 + c), mean_in[k], acc);
    red[sl][j] = acc;
    {   // P' rows of this CTA: warp w takes rows 8w .. 8w+7, lane l the 8 columns 8l .. 8l+7 of both planes
        const int warp = tid >> 5, lane = tid & 31;
        const float4 m0 = *reinterpret_cast<const float4 *>(mean_in + 256 + lane * 8), m1 = *reinterpret_cast<const float4 *>(mean_in + 260 + lane * 8);
        const float m[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {                         // unrolled: the 16 loads of a warp's rows are in flight together
            const long long off = (((long long)(p_batched ? b : 0) * 512) + r * 64 + warp * 8 + i) * 256 + lane * 8;
            const uint4 h4 = __ldg(reinterpret_cast<const uint4 *>(Ph + off)), l4 = __ldg(reinterpret_cast<const uint4 *>(Pl + off));
            const __half2 *hh = reinterpret_cast<const __half2 *>(&h4), *ll = reinterpret_cast<const __half2 *>(&l4);
            float d = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 a = __half22float2(hh[q]), e = __half22float2(ll[q]);
                d = fmaf(a.x + e.x, m[2 * q], d); d = fmaf(a.y + e.y, m[2 * q + 1], d);
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
            if (lane == 0) pdot[warp * 8 + i] = d * (1.f / lintc::kScale);
        }
    }
    __syncthreads();
    if (tid < 64) bias_out[(long long)b * 512 + r * 64 + tid] = -(((red[0][tid] + red[1][tid]) + (red[2][tid] + red[3][tid])) + pdot[tid]);
}

// InstanceNorm statistics of the centred layer-0 output (OPB_LIN_CSTATS_RELU partials) -> rstd, folded straight into the
// per-frame layer-1 weights: W1s[b][o][c] = W1[o][c] * rstd[b][c] as split-precision planes.  CTA (b, g) owns channels
// [64g, 64g+64): it needs nobody else's statistics.
__global__ void __launch_bounds__(256)
k_instnorm_scale_weight_planes(const float2 *__restrict__ part, int parts, int rows, const int *__restrict__ row_counts, float eps,
                               const float *__restrict__ W, int Cout, int K, __half *__restrict__ hi, __half *__restrict__ lo,
                               float *__restrict__ rstd_out) {
    __shared__ double red_s[16][64], red_q[16][64];
    __shared__ float rs[64];
    const int b = blockIdx.x, g = blockIdx.y, tid = threadIdx.x;
    if (row_counts) rows = max(1, min(rows, __ldg(row_counts + b)));      // ragged batch: statistics over this frame's valid rows
    {   // thread = (slice of 16, 4 channels): 32-byte loads, many in flight (four slices of scalar loads were a chain of L2 round trips)
        const int fs = tid >> 4, f4 = (tid & 15) * 4;
        double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
        const float4 *p4 = reinterpret_cast<const float4 *>(part + (long long)b * parts * K + g * 64 + f4);
#pragma unroll 4
        for (int t = fs; t < parts; t += 16) {
            const float4 a = __ldg(p4 + (long long)t * (K / 2)), c = __ldg(p4 + (long long)t * (K / 2) + 1);
            s[0] += (double)a.x; q[0] += (double)a.y; s[1] += (double)a.z; q[1] += (double)a.w;
            s[2] += (double)c.x; q[2] += (double)c.y; s[3] += (double)c.z; q[3] += (double)c.w;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) { red_s[fs][f4 + i] = s[i]; red_q[fs][f4 + i] = q[i]; }
    }
    __syncthreads();
    if (tid < 64) {
        double ss = 0, qq = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) { ss += red_s[k][tid]; qq += red_q[k][tid]; }
        const double mu = ss / rows;
        double var = qq / rows - mu * mu;
        if (var < 0) var = 0;
        rs[tid] = (float)(1.0 / sqrt(var + (double)eps));
        if (rstd_out) rstd_out[(long long)b * K + g * 64 + tid] = rs[tid];
    }
    __syncthreads();
    const int c4 = (tid & 15) * 4;                           // 16 threads per weight row segment of 64 channels
    const float4 r4 = *reinterpret_cast<const float4 *>(rs + c4);
    for (int o = tid >> 4; o < Cout; o += 16) {
        const float4 w = __ldg(reinterpret_cast<const float4 *>(W + (long long)o * K + g * 64 + c4));
        __half h[4], l[4];
        tc::split_f16(w.x * r4.x * lintc::kScale, h[0], l[0]); tc::split_f16(w.y * r4.y * lintc::kScale, h[1], l[1]);
        tc::split_f16(w.z * r4.z * lintc::kScale, h[2], l[2]); tc::split_f16(w.w * r4.w * lintc::kScale, h[3], l[3]);
        const long long off = ((long long)b * Cout + o) * K + g * 64 + c4;
        *reinterpret_cast<uint2 *>(hi + off) = *reinterpret_cast<uint2 *>(h);
        *reinterpret_cast<uint2 *>(lo + off) = *reinterpret_cast<uint2 *>(l);
    }
}

// W1s[b][o][c] = W1[o][c] * rstd[b][c] as split-precision planes: the 1/std of InstanceNorm folded into the next layer
__global__ void k_scale_weight_planes(const float *__restrict__ W, const float *__restrict__ rstd, int Cout, int K,
                                      __half *__restrict__ hi, __half *__restrict__ lo) {
    const int b = blockIdx.y;
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= (long long)Cout * K) return;
    const int c = (int)(i % K);
    const float4 w = *reinterpret_cast<const float4 *>(W + i);
    const float4 r = *reinterpret_cast<const float4 *>(rstd + (long long)b * K + c);
    __half h[4], l[4];
    tc::split_f16(w.x * r.x * lintc::kScale, h[0], l[0]); tc::split_f16(w.y * r.y * lintc::kScale, h[1], l[1]);
    tc::split_f16(w.z * r.z * lintc::kScale, h[2], l[2]); tc::split_f16(w.w * r.w * lintc::kScale, h[3], l[3]);
    const long long o = (long long)b * Cout * K + i;
    *reinterpret_cast<uint2 *>(hi + o) = *reinterpret_cast<uint2 *>(h);
    *reinterpret_cast<uint2 *>(lo + o) = *reinterpret_cast<uint2 *>(l);
}

static int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

}  // namespace opb

extern "C" int op_split_planes(const float *x, void *hi, void *lo, long long n, void *stream) {
    OPB_CHECK_ARG(x && hi && lo && n >= 0 && n % 4 == 0, "op_split_planes: bad argument (n=%lld must be a multiple of 4)", n);
    if (n == 0) return OPB_OK;
    opb::k_split_planes64<<<opb::ceil_div(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(x, (__half *)hi, (__half *)lo, n);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_norm_relu_split(const float *h, const float *mean, const float *rstd, void *hi, void *lo, int B, int rows,
                                  int C, void *stream) {
    OPB_CHECK_ARG(h && mean && rstd && hi && lo, "op_norm_relu_split: null pointer");
    OPB_CHECK_ARG(B >= 0 && rows >= 0 && C > 0 && C % 4 == 0, "op_norm_relu_split: bad shape B=%d rows=%d C=%d", B, rows, C);
    const long long n = (long long)B * rows * C;
    if (n == 0) return OPB_OK;
    opb::k_norm_relu_split<<<opb::ceil_div(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(h, mean, rstd, (__half *)hi, (__half *)lo,
                                                                                      (long long)rows * C, C, n);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_colsum_partial(const float *x, int B, int rows, int C, float *partial, void *stream) {
    OPB_CHECK_ARG(x && partial && B > 0 && rows > 0 && C > 0 && B <= 65535, "op_colsum_partial: bad argument");
    const int parts = opb::ceil_div(rows, opb::lintc::BM) * 4;
    opb::k_colsum_partial<<<dim3(parts, B), 256, 0, (cudaStream_t)stream>>>(x, rows, C, parts, (float2 *)partial);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_mlp_center_bias_ragged(const float *xpart, int xparts, const float *qpart, int qparts, const float *W0aT, const void *Ph,
                                         const void *Pl, int p_batched, int B, int rows, const int *row_counts, float *bias_out, void *stream) {
    OPB_CHECK_ARG(xpart && qpart && W0aT && Ph && Pl && bias_out && B > 0 && rows > 0 && xparts > 0 && qparts > 0 && B <= 65535,
                  "op_mlp_center_bias: bad argument");
    opb::k_mlp_center_bias<<<dim3(B, 8), 256, 0, (cudaStream_t)stream>>>((const float2 *)xpart, xparts, (const float2 *)qpart, qparts, W0aT,
                                                                         (const __half *)Ph, (const __half *)Pl, p_batched, rows, row_counts, bias_out);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_mlp_center_bias(const float *xpart, int xparts, const float *qpart, int qparts, const float *W0aT, const void *Ph,
                                  const void *Pl, int p_batched, int B, int rows, float *bias_out, void *stream) {
    return op_mlp_center_bias_ragged(xpart, xparts, qpart, qparts, W0aT, Ph, Pl, p_batched, B, rows, nullptr, bias_out, stream);
}

extern "C" int op_scale_weight_planes(const float *W, const float *rstd, int B, int Cout, int K, void *hi, void *lo, void *stream) {
    OPB_CHECK_ARG(W && rstd && hi && lo && B > 0 && Cout > 0 && K > 0 && K % 4 == 0 && B <= 65535, "op_scale_weight_planes: bad argument");
    opb::k_scale_weight_planes<<<dim3(opb::ceil_div((long long)Cout * K / 4, 256), B), 256, 0, (cudaStream_t)stream>>>(W, rstd, Cout, K, (__half *)hi, (__half *)lo);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_instnorm_scale_weight_planes_ragged(const float *stats_partial, int B, int parts, int rows_per_batch, const int *row_counts,
                                                      float eps, const float *W, int Cout, int K, void *hi, void *lo, float *rstd, void *stream) {
    OPB_CHECK_ARG(stats_partial && W && hi && lo, "op_instnorm_scale_weight_planes: null pointer");
    OPB_CHECK_ARG(B > 0 && B <= 65535 && parts > 0 && rows_per_batch > 0 && Cout > 0 && K > 0 && K % 64 == 0,
                  "op_instnorm_scale_weight_planes: bad shape B=%d parts=%d rows=%d Cout=%d K=%d", B, parts, rows_per_batch, Cout, K);
    opb::k_instnorm_scale_weight_planes<<<dim3(B, K / 64), 256, 0, (cudaStream_t)stream>>>((const float2 *)stats_partial, parts, rows_per_batch, row_counts,
                                                                                           eps, W, Cout, K, (__half *)hi, (__half *)lo, rstd);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_instnorm_scale_weight_planes(const float *stats_partial, int B, int parts, int rows_per_batch, float eps, const float *W,
                                               int Cout, int K, void *hi, void *lo, float *rstd, void *stream) {
    return op_instnorm_scale_weight_planes_ragged(stats_partial, B, parts, rows_per_batch, nullptr, eps, W, Cout, K, hi, lo, rstd, stream);
}

extern "C" int op_linear_tc_stat_parts(int rows_per_batch) { return opb::ceil_div(rows_per_batch, opb::lintc::BM) * 4; }

extern "C" int op_instnorm_finalize_tc(const float *stats_partial, const float *bias, int B, int parts, int C, int rows_per_batch,
                                       float eps, float *mean, float *rstd, void *stream) {
    OPB_CHECK_ARG(stats_partial && mean && rstd, "op_instnorm_finalize_tc: null pointer");
    OPB_CHECK_ARG(B > 0 && parts > 0 && C > 0 && rows_per_batch > 0, "op_instnorm_finalize_tc: bad shape");
    dim3 grid(opb::ceil_div(C, 128), B);
    opb::k_instnorm_finalize_tc<<<grid, 128, 0, (cudaStream_t)stream>>>((const float2 *)stats_partial, bias, parts, C, rows_per_batch,
                                                                         eps, mean, rstd);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_linear_tc_ragged(const void *A1h, const void *A1l, int K1, const void *A2h, const void *A2l, int K2, const void *Wh,
                                   const void *Wl, const void *W2h, const void *W2l, const float *bias, float *C, void *Ch, void *Cl, int B, int rows_per_batch, int Cout,
                                   int flags, int elu_cols, float *stats_partial, const float *ksum, float zscale_num, const int *row_counts, void *stream) {
    using namespace opb;
    OPB_CHECK_ARG(!row_counts || !(flags & OPB_LIN_STATS), "op_linear_tc_ragged: OPB_LIN_STATS is not available with row_counts (use OPB_LIN_CSTATS_RELU)");
    OPB_CHECK_ARG(A1h && A1l && Wh && Wl, "op_linear_tc: null operand plane");
    OPB_CHECK_ARG(K1 > 0 && K1 % lintc::BK == 0 && K2 >= 0 && K2 % lintc::BK == 0, "op_linear_tc: K1=%d K2=%d must be multiples of 64", K1, K2);
    OPB_CHECK_ARG((K2 == 0) == (A2h == nullptr) && (A2h == nullptr) == (A2l == nullptr), "op_linear_tc: A2/K2 mismatch");
    OPB_CHECK_ARG((W2h == nullptr) == (W2l == nullptr) && (!W2h || K2), "op_linear_tc: W2 planes come together and need A2");
    OPB_CHECK_ARG(!(flags & OPB_LIN_W2BATCHED) || W2h, "op_linear_tc: W2BATCHED without W2");
    OPB_CHECK_ARG(Cout > 0 && Cout % lintc::BN == 0 && Cout <= lintc::BIAS_FLOATS, "op_linear_tc: Cout=%d must be 256 or 512", Cout);
    OPB_CHECK_ARG(C || Ch, "op_linear_tc: no output requested");
    OPB_CHECK_ARG((Ch == nullptr) == (Cl == nullptr), "op_linear_tc: Ch/Cl must come together");
    OPB_CHECK_ARG(!(flags & OPB_LIN_RESIDUAL) || C || Ch, "op_linear_tc: RESIDUAL needs an output to add to");
    {
        const int nst = ((flags & OPB_LIN_STATS) != 0) + ((flags & OPB_LIN_COLSUM) != 0) + ((flags & OPB_LIN_CSTATS_RELU) != 0);
        OPB_CHECK_ARG(nst == 0 || stats_partial, "op_linear_tc: STATS / COLSUM / CSTATS_RELU need stats_partial");
        OPB_CHECK_ARG(nst <= 1, "op_linear_tc: STATS, COLSUM and CSTATS_RELU are exclusive");
        OPB_CHECK_ARG(!(flags & OPB_LIN_BIASB) || bias, "op_linear_tc: BIASB needs a bias");
    }
    OPB_CHECK_ARG(!(flags & OPB_LIN_ZSCALE) || (ksum && Cout == 256), "op_linear_tc: ZSCALE needs ksum and Cout == 256");
    OPB_CHECK_ARG(!(flags & OPB_LIN_NORM_RELU), "op_linear_tc: NORM_RELU is applied by op_norm_relu_split");
    OPB_CHECK_ARG(!(flags & OPB_LIN_ELU1) || elu_cols % 32 == 0, "op_linear_tc: elu_cols must be a multiple of 32");
    OPB_CHECK_ARG(B >= 0 && rows_per_batch >= 0, "op_linear_tc: bad B=%d rows=%d", B, rows_per_batch);
    const bool unit8 = (flags & OPB_LIN_UNIT_PLANES8) != 0;
    OPB_CHECK_ARG(!unit8 || (Cout == 256 && Ch && Cl && !C && !(flags & (lintc::kEpiMask & ~OPB_LIN_UNIT_PLANES8))),
                  "op_linear_tc: UNIT_PLANES8 needs Cout == 256, plane outputs only and no other epilogue flag");
    if (B == 0 || rows_per_batch == 0) return OPB_OK;
    const int K = W2h ? K1 : K1 + K2;                 // columns of W
    CUtensorMap ta1h, ta1l, ta2h, ta2l, twh, twl, tw2h, tw2l, tc_, tch, tcl;
    int rc;
    // CTA pairs (cta_group::2) by default; OPB_LINEAR_CG=1 selects the one-CTA-per-tile form of the same kernel (A/B runs)
    static const int cg = [] { const char *e = getenv("OPB_LINEAR_CG"); return (e && e[0] == '1') ? 1 : 2; }();
    const int wbox = lintc::BN / cg;                  // rows of a W block one CTA loads
    const int a1b = (flags & OPB_LIN_A1_BCAST) ? 1 : B, a2b = (flags & OPB_LIN_A2_BCAST) ? 1 : B;
    const bool w2_ident = (flags & OPB_LIN_W2_IDENTITY) && cg == 2;
    OPB_CHECK_ARG(!(flags & OPB_LIN_W2_IDENTITY) || (W2h && K2 == 256 && Cout == 256 && !(flags & OPB_LIN_W2BATCHED)),
                  "op_linear_tc: W2_IDENTITY needs unbatched W2 planes with K2 = Cout = 256");
    // OPB_LIN_BK=32 (A/B runs): k-blocks of 32 in six ring slots (Cfg<2, 32>) for the specialised instantiations without the identity
    // residual -- the same 192 KB ring in finer slots, tried against the ring's round trip.  Measured slower (profiles/r02i_linear_bk32.txt:
    // K' 221 -> 229 us, MLP0 668 -> 725 us; 23.3 -> 24.4 ms per bench step): twice the barrier / TMA traffic and 64-byte rows cost more
    // than the two extra slots in flight bring.  The operand maps follow the kernel's k-block (box columns, swizzle).
    static const bool want_k32 = [] { const char *e = getenv("OPB_LIN_BK"); return e && !strcmp(e, "32"); }();
    static const int dbg = [] { const char *e = getenv("OPB_LIN_DBG"); return e ? atoi(e) : 0; }();
    const int key = (flags & lintc::kEpiMask) | (C ? lintc::kHasC : 0) | (Ch ? lintc::kHasP : 0);
    constexpr int kK = OPB_LIN_ELU1 | OPB_LIN_COLSUM | lintc::kHasP;                          // K' projection
    constexpr int kQ = OPB_LIN_ELU1 | OPB_LIN_COLSUM | OPB_LIN_ZSCALE | lintc::kHasP;         // Q projection
    constexpr int kM0 = OPB_LIN_BIASB | OPB_LIN_CSTATS_RELU | lintc::kHasP;                   // MLP layer 0
    constexpr int kM1 = OPB_LIN_COLSUM | lintc::kHasP;                                        // MLP layer 1 (residual inside the GEMM)
    constexpr int kF = lintc::kHasC;                                                          // final projection (fp32 output)
    constexpr int kFN = OPB_LIN_UNIT_PLANES8 | lintc::kHasP;                                  // final projection + F.normalize -> similarity operand planes
    const bool k32 = want_k32 && cg == 2 && !dbg && !w2_ident && (key == kK || key == kQ || key == kM0 || key == kF || key == kFN);
    auto amap = [&](CUtensorMap *m, const void *base, int batch, int rows, int cols, int box_rows) {
        return k32 ? make_tmap_f16_3d_k32(m, base, batch, rows, cols, box_rows) : make_tmap_f16_3d(m, base, batch, rows, cols, box_rows);
    };
    if ((rc = amap(&ta1h, A1h, a1b, rows_per_batch, K1, lintc::BM)) || (rc = amap(&ta1l, A1l, a1b, rows_per_batch, K1, lintc::BM)) ||
        (rc = amap(&twh, Wh, (flags & OPB_LIN_WBATCHED) ? B : 1, Cout, K, wbox)) ||
        (rc = amap(&twl, Wl, (flags & OPB_LIN_WBATCHED) ? B : 1, Cout, K, wbox)))
        return rc;
    if (K2) {
        if ((rc = amap(&ta2h, A2h, a2b, rows_per_batch, K2, lintc::BM)) || (rc = amap(&ta2l, A2l, a2b, rows_per_batch, K2, lintc::BM)))
            return rc;
    } else {
        ta2h = ta1h; ta2l = ta1l;
    }
    if (W2h) {
        const int wb = (flags & OPB_LIN_W2BATCHED) ? B : 1;
        if ((rc = amap(&tw2h, W2h, wb, Cout, K2, w2_ident ? 32 : wbox)) || (rc = amap(&tw2l, W2l, wb, Cout, K2, wbox))) return rc;
    } else {
        tw2h = twh; tw2l = twl;
    }
    OPB_CHECK_ARG(((uintptr_t)C | (uintptr_t)Ch | (uintptr_t)Cl) % 16 == 0, "op_linear_tc: outputs must be 16-byte aligned");
    if (C) { if ((rc = make_tmap_f32_3d_store(&tc_, C, B, rows_per_batch, Cout, 32))) return rc; } else tc_ = ta1h;
    if (unit8) {
        // similarity operand planes: hi fp16, then two byte planes in the Cl buffer (lo e5m2 | hi e4m3) -- the byte maps ride in the
        // kernel's tm_cl / tm_c parameters
        uint8_t *bytes = (uint8_t *)Cl;
        if ((rc = make_tmap_f16_3d_store(&tch, Ch, B, rows_per_batch, Cout, 32)) || (rc = make_tmap_u8_3d_store(&tcl, bytes, B, rows_per_batch, Cout, 32)) ||
            (rc = make_tmap_u8_3d_store(&tc_, bytes + (size_t)B * rows_per_batch * Cout, B, rows_per_batch, Cout, 32)))
            return rc;
    } else if (Ch) {
        if ((rc = make_tmap_f16_3d_store(&tch, Ch, B, rows_per_batch, Cout, 32)) || (rc = make_tmap_f16_3d_store(&tcl, Cl, B, rows_per_batch, Cout, 32))) return rc;
    } else { tch = ta1h; tcl = ta1h; }
    lintc::Params p{};
    p.B = B; p.rows = rows_per_batch; p.Cout = Cout; p.K1 = K1; p.K2 = K2; p.flags = flags; p.elu_cols = elu_cols;
    p.row_tiles = ceil_div(rows_per_batch, lintc::BM); p.col_tiles = Cout / lintc::BN;
    p.row_groups = ceil_div(p.row_tiles, cg);
    const long long total = (long long)B * p.row_groups * p.col_tiles;
    OPB_CHECK_ARG(total < (1ll << 31), "op_linear_tc: too many tiles");
    p.total_tiles = (int)total;
    p.bias = bias; p.C = C; p.Ch = (__half *)Ch; p.Cl = (__half *)Cl; p.stats = (float2 *)stats_partial;
    p.ksum = ksum; p.zscale_num = zscale_num; p.w_batched = (flags & OPB_LIN_WBATCHED) ? 1 : 0;
    p.w2 = W2h ? ((flags & OPB_LIN_W2BATCHED) ? 2 : 1) : 0;
    p.w2_ident = w2_ident ? 1 : 0;
    p.a_bcast = ((flags & OPB_LIN_A1_BCAST) ? 1 : 0) | ((flags & OPB_LIN_A2_BCAST) ? 2 : 0);
    p.row_counts = row_counts;
    // L2 prefetch of the A boxes ahead of their TMA loads: measured SLOWER at every distance (profiles/r02h_linear_prefetch.txt:
    // K' 224 -> 230 / 233 / 289 us, MLP0 695 -> 726 / 727 / 732 us at 2 / 4 / 8 k-blocks ahead), so it is off; the switch stays for A/B runs
    static const int prefetch = [] { const char *e = getenv("OPB_LIN_PF"); return e ? atoi(e) : 0; }();
    p.prefetch = prefetch;
    p.dbg = dbg;
    const int groups = sm_count() / cg;                                   // persistent CTAs (or CTA pairs), one per SM
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    cfg.gridDim = dim3((unsigned)((p.total_tiles < groups ? p.total_tiles : groups) * cg));
    cfg.blockDim = dim3(lintc::THREADS);
    cfg.dynamicSmemBytes = cg == 2 ? lintc::Cfg<2>::SMEM_BYTES : lintc::Cfg<1>::SMEM_BYTES;
    cfg.stream = (cudaStream_t)stream;
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cg; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    // one specialised instantiation per layer kind of the matcher (epilogue flags and output kinds as compile-time constants), the
    // generic kernel for everything else, for the one-CTA form and whenever experiment switches are set
    auto launch = [&](auto kern, int smem_bytes) -> cudaError_t {
        static const void *configured[24];       // kernels whose dynamic shared memory limit has been raised (all share one pointer type)
        static int n_configured = 0;
        bool seen = false;
        for (int i = 0; i < n_configured; ++i) seen |= configured[i] == (const void *)kern;
        if (!seen) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
            if (e != cudaSuccess) return e;
            if (n_configured < 24) configured[n_configured++] = (const void *)kern;
        }
        return cudaLaunchKernelEx(&cfg, kern, ta1h, ta1l, ta2h, ta2l, twh, twl, tw2h, tw2l, tc_, tch, tcl, p);
    };
    const int smem2 = lintc::Cfg<2>::SMEM_BYTES;
    static_assert(lintc::Cfg<2, 32>::SMEM_BYTES == lintc::Cfg<2>::SMEM_BYTES, "both ring forms fill the same shared memory");
    OPB_CHECK_ARG(!unit8 || (cg == 2 && !dbg), "op_linear_tc: UNIT_PLANES8 exists in the pair form only (unset OPB_LINEAR_CG / OPB_LIN_DBG)");
    if (cg == 1) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<1, -1>, lintc::Cfg<1>::SMEM_BYTES));
    else if (dbg) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, -1>, smem2));
    else if (key == kK && k32) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kK, 32>, smem2));
    else if (key == kQ && k32) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kQ, 32>, smem2));
    else if (key == kM0 && k32) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kM0, 32>, smem2));
    else if (key == kF && k32) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kF, 32>, smem2));
    else if (key == kFN && k32) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kFN, 32>, smem2));
    else if (key == kK) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kK>, smem2));
    else if (key == kQ) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kQ>, smem2));
    else if (key == kM0) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kM0>, smem2));
    else if (key == kM1) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kM1>, smem2));
    else if (key == kF) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kF>, smem2));
    else if (key == kFN) OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, kFN>, smem2));
    else OPB_CHECK_CUDA(launch(lintc::k_linear_tc<2, -1>, smem2));
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_linear_tc(const void *A1h, const void *A1l, int K1, const void *A2h, const void *A2l, int K2, const void *Wh,
                            const void *Wl, const void *W2h, const void *W2l, const float *bias, float *C, void *Ch, void *Cl, int B, int rows_per_batch, int Cout,
                            int flags, int elu_cols, float *stats_partial, const float *ksum, float zscale_num, void *stream) {
    return op_linear_tc_ragged(A1h, A1l, K1, A2h, A2l, K2, Wh, Wl, W2h, W2l, bias, C, Ch, Cl, B, rows_per_batch, Cout, flags, elu_cols, stats_partial, ksum,
                               zscale_num, nullptr, stream);
}
