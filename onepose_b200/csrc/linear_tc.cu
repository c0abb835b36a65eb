// (2) Fused linear layer on the 5th-gen tensor cores -- the 1x1 Conv1d projections of MultiHeadedAttention / MLP /
// final_proj (GATs_SuperGlue.py:108-120,135-147,194-199,241).  Kernel in linear_tc.cuh; this file holds the
// elementwise producers of the split-precision planes and the C ABI.
#include "common.cuh"
#include "linear_tc.cuh"

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

extern "C" int op_linear_tc(const void *A1h, const void *A1l, int K1, const void *A2h, const void *A2l, int K2, const void *Wh,
                            const void *Wl, const float *bias, float *C, void *Ch, void *Cl, int B, int rows_per_batch, int Cout,
                            int flags, int elu_cols, float *stats_partial, const float *ksum, float zscale_num, void *stream) {
    using namespace opb;
    OPB_CHECK_ARG(A1h && A1l && Wh && Wl, "op_linear_tc: null operand plane");
    OPB_CHECK_ARG(K1 > 0 && K1 % lintc::BK == 0 && K2 >= 0 && K2 % lintc::BK == 0, "op_linear_tc: K1=%d K2=%d must be multiples of 64", K1, K2);
    OPB_CHECK_ARG((K2 == 0) == (A2h == nullptr) && (A2h == nullptr) == (A2l == nullptr), "op_linear_tc: A2/K2 mismatch");
    OPB_CHECK_ARG(Cout > 0 && Cout % lintc::BN == 0 && Cout <= lintc::BIAS_FLOATS, "op_linear_tc: Cout=%d must be 256 or 512", Cout);
    OPB_CHECK_ARG(C || Ch, "op_linear_tc: no output requested");
    OPB_CHECK_ARG((Ch == nullptr) == (Cl == nullptr), "op_linear_tc: Ch/Cl must come together");
    OPB_CHECK_ARG(!(flags & OPB_LIN_RESIDUAL) || C, "op_linear_tc: RESIDUAL needs the fp32 output");
    OPB_CHECK_ARG(!(flags & (OPB_LIN_STATS | OPB_LIN_COLSUM)) || stats_partial, "op_linear_tc: STATS / COLSUM need stats_partial");
    OPB_CHECK_ARG(!((flags & OPB_LIN_STATS) && (flags & OPB_LIN_COLSUM)), "op_linear_tc: STATS and COLSUM are exclusive");
    OPB_CHECK_ARG(!(flags & OPB_LIN_ZSCALE) || (ksum && Cout == 256), "op_linear_tc: ZSCALE needs ksum and Cout == 256");
    OPB_CHECK_ARG(!(flags & OPB_LIN_NORM_RELU), "op_linear_tc: NORM_RELU is applied by op_norm_relu_split");
    OPB_CHECK_ARG(!(flags & OPB_LIN_ELU1) || elu_cols % 32 == 0, "op_linear_tc: elu_cols must be a multiple of 32");
    OPB_CHECK_ARG(B >= 0 && rows_per_batch >= 0, "op_linear_tc: bad B=%d rows=%d", B, rows_per_batch);
    if (B == 0 || rows_per_batch == 0) return OPB_OK;
    const int K = K1 + K2;
    CUtensorMap ta1h, ta1l, ta2h, ta2l, twh, twl, tc_, tch, tcl;
    int rc;
    if ((rc = make_tmap_f16_3d(&ta1h, A1h, B, rows_per_batch, K1, lintc::BM)) || (rc = make_tmap_f16_3d(&ta1l, A1l, B, rows_per_batch, K1, lintc::BM)) ||
        (rc = make_tmap_f16_3d(&twh, Wh, (flags & OPB_LIN_WBATCHED) ? B : 1, Cout, K, lintc::BN)) ||
        (rc = make_tmap_f16_3d(&twl, Wl, (flags & OPB_LIN_WBATCHED) ? B : 1, Cout, K, lintc::BN)))
        return rc;
    if (K2) {
        if ((rc = make_tmap_f16_3d(&ta2h, A2h, B, rows_per_batch, K2, lintc::BM)) || (rc = make_tmap_f16_3d(&ta2l, A2l, B, rows_per_batch, K2, lintc::BM)))
            return rc;
    } else {
        ta2h = ta1h; ta2l = ta1l;
    }
    OPB_CHECK_ARG(((uintptr_t)C | (uintptr_t)Ch | (uintptr_t)Cl) % 16 == 0, "op_linear_tc: outputs must be 16-byte aligned");
    if (C) { if ((rc = make_tmap_f32_3d_store(&tc_, C, B, rows_per_batch, Cout, 32))) return rc; } else tc_ = ta1h;
    if (Ch) {
        if ((rc = make_tmap_f16_3d_store(&tch, Ch, B, rows_per_batch, Cout, 32)) || (rc = make_tmap_f16_3d_store(&tcl, Cl, B, rows_per_batch, Cout, 32))) return rc;
    } else { tch = ta1h; tcl = ta1h; }
    lintc::Params p{};
    p.B = B; p.rows = rows_per_batch; p.Cout = Cout; p.K1 = K1; p.K2 = K2; p.flags = flags; p.elu_cols = elu_cols;
    p.row_tiles = ceil_div(rows_per_batch, lintc::BM); p.col_tiles = Cout / lintc::BN;
    const long long total = (long long)B * p.row_tiles * p.col_tiles;
    OPB_CHECK_ARG(total < (1ll << 31), "op_linear_tc: too many tiles");
    p.total_tiles = (int)total;
    p.bias = bias; p.C = C; p.Ch = (__half *)Ch; p.Cl = (__half *)Cl; p.stats = (float2 *)stats_partial;
    p.ksum = ksum; p.zscale_num = zscale_num; p.w_batched = (flags & OPB_LIN_WBATCHED) ? 1 : 0;
    static bool configured = false;
    if (!configured) {
        OPB_CHECK_CUDA(cudaFuncSetAttribute(lintc::k_linear_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, lintc::SMEM_BYTES));
        configured = true;
    }
    const int grid = p.total_tiles < sm_count() ? p.total_tiles : sm_count();
    lintc::k_linear_tc<<<grid, lintc::THREADS, lintc::SMEM_BYTES, (cudaStream_t)stream>>>(ta1h, ta1l, ta2h, ta2l, twh, twl, tc_, tch, tcl, p);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
