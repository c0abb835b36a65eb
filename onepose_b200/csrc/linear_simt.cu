// (2) Fused linear layer, fp32 CUDA-core version -- the 1x1 Conv1d projections of
// MultiHeadedAttention / MLP / final_proj (GATs_SuperGlue.py:108-120,135-147,194-199,241).
//
// C[b,n,:] = epi([A1 | A2][b,n,:] . W^T + bias).  128x128x16 tiles, 256 threads, 8x8 register
// micro-tiles, register-prefetch software pipeline.  The epilogue fuses what the reference runs as
// separate ops: elu+1 feature map, residual add, InstanceNorm statistics (float64 partial sums,
// ordered -> deterministic), and the A-side InstanceNorm+ReLU of the second MLP layer.
// This is the exact-fp32 path; the tcgen05 split-precision version lives in linear_tc.cu.
#include "common.cuh"

namespace opb {

constexpr int LBM = 128, LBN = 128, LBK = 16, LPAD = 4;

struct LinearParams {
    const float *A1; int K1;
    const float *A2; int K2;
    const float *W; const float *bias;
    float *C; int rows; int Cout; int flags; int elu_cols;
    const float *mean; const float *rstd; double *stats; int row_tiles;
};

__global__ void __launch_bounds__(256)
k_linear_simt(const LinearParams p) {
    __shared__ __align__(16) float As[LBK][LBM + LPAD];
    __shared__ __align__(16) float Ws[LBK][LBN + LPAD];
    __shared__ double red[2][8][LBN];

    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int b = blockIdx.z, m0 = blockIdx.x * LBM, n0 = blockIdx.y * LBN;
    const int K = p.K1 + p.K2;
    const long long rowbase = (long long)b * p.rows;

    // global->register staging: each thread moves 2 float4 of A and 2 of W per k-tile
    const int lr = tid >> 2, lk = (tid & 3) * 4;      // rows lr and lr+64, k offset lk
    float4 ra[2], rw[2];
    auto load_tile = [&](int k0) {
        const float *Ab; int kk, ld;
        if (k0 < p.K1) { Ab = p.A1; kk = k0; ld = p.K1; } else { Ab = p.A2; kk = k0 - p.K1; ld = p.K2; }
        const bool norm = (p.flags & OPB_LIN_NORM_RELU) && k0 < p.K1;
        float4 mu = make_float4(0, 0, 0, 0), rs = make_float4(1, 1, 1, 1);
        if (norm) {
            mu = *reinterpret_cast<const float4 *>(p.mean + (long long)b * p.K1 + kk + lk);
            rs = *reinterpret_cast<const float4 *>(p.rstd + (long long)b * p.K1 + kk + lk);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = m0 + lr + h * 64;
            float4 v = make_float4(0, 0, 0, 0);
            if (r < p.rows) v = *reinterpret_cast<const float4 *>(Ab + (rowbase + r) * ld + kk + lk);
            if (norm) {
                v.x = fmaxf((v.x - mu.x) * rs.x, 0.f); v.y = fmaxf((v.y - mu.y) * rs.y, 0.f);
                v.z = fmaxf((v.z - mu.z) * rs.z, 0.f); v.w = fmaxf((v.w - mu.w) * rs.w, 0.f);
            }
            ra[h] = v;
            rw[h] = *reinterpret_cast<const float4 *>(p.W + (long long)(n0 + lr + h * 64) * K + k0 + lk);
        }
    };
    auto store_tile = [&]() {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = lr + h * 64;
            As[lk + 0][r] = ra[h].x; As[lk + 1][r] = ra[h].y; As[lk + 2][r] = ra[h].z; As[lk + 3][r] = ra[h].w;
            Ws[lk + 0][r] = rw[h].x; Ws[lk + 1][r] = rw[h].y; Ws[lk + 2][r] = rw[h].z; Ws[lk + 3][r] = rw[h].w;
        }
    };

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    load_tile(0);
    for (int k0 = 0; k0 < K; k0 += LBK) {
        store_tile();
        __syncthreads();
        if (k0 + LBK < K) load_tile(k0 + LBK);
#pragma unroll
        for (int k = 0; k < LBK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[k][64 + ty * 4]);
            const float4 w0 = *reinterpret_cast<const float4 *>(&Ws[k][tx * 4]);
            const float4 w1 = *reinterpret_cast<const float4 *>(&Ws[k][64 + tx * 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
        }
        __syncthreads();
    }

    // ---- epilogue ----
    float bias[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bias[j] = p.bias ? p.bias[n0 + (j >> 2) * 64 + tx * 4 + (j & 3)] : 0.f;
    double csum[8], csq[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) csum[j] = csq[j] = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = m0 + (i >> 2) * 64 + ty * 4 + (i & 3);
        const bool valid = r < p.rows;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
            float v = acc[i][j] + bias[j];
            if ((p.flags & OPB_LIN_ELU1) && c < p.elu_cols) v = elu1(v);
            acc[i][j] = v;
            if ((p.flags & OPB_LIN_STATS) && valid) { csum[j] += (double)v; csq[j] += (double)v * (double)v; }
        }
        if (valid) {
            float *crow = p.C + (rowbase + r) * p.Cout + n0 + tx * 4;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float4 v = make_float4(acc[i][h * 4], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]);
                float4 *dst = reinterpret_cast<float4 *>(crow + h * 64);
                if (p.flags & OPB_LIN_RESIDUAL) { const float4 o = *dst; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
                *dst = v;
            }
        }
    }
    if (p.flags & OPB_LIN_STATS) {
        // rows of a warp: ty in {2w, 2w+1}; fold the pair with a shuffle, then 8 warps through smem in fixed order
        const int warp = tid >> 5;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            csum[j] += __shfl_xor_sync(0xffffffffu, csum[j], 16);
            csq[j] += __shfl_xor_sync(0xffffffffu, csq[j], 16);
        }
        if ((tid & 16) == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int cl = (j >> 2) * 64 + tx * 4 + (j & 3);
                red[0][warp][cl] = csum[j];
                red[1][warp][cl] = csq[j];
            }
        }
        __syncthreads();
        if (tid < LBN) {
            double s = 0, q = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) { s += red[0][w][tid]; q += red[1][w][tid]; }
            double *dst = p.stats + (((long long)b * p.row_tiles + blockIdx.x) * p.Cout + n0 + tid) * 2;
            dst[0] = s; dst[1] = q;
        }
    }
}

// mean / rstd from ordered float64 partials
__global__ void k_instnorm_finalize(const double *__restrict__ part, int row_tiles, int C, int rows, float eps,
                                    float *__restrict__ mean, float *__restrict__ rstd) {
    const int b = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s = 0, q = 0;
    for (int t = 0; t < row_tiles; ++t) {
        const double *src = part + (((long long)b * row_tiles + t) * C + c) * 2;
        s += src[0]; q += src[1];
    }
    const double mu = s / rows;
    double var = q / rows - mu * mu;     // biased variance, InstanceNorm1d default (GATs_SuperGlue.py:144)
    if (var < 0) var = 0;
    mean[(long long)b * C + c] = (float)mu;
    rstd[(long long)b * C + c] = (float)(1.0 / sqrt(var + (double)eps));
}

__global__ void k_l2_normalize_rows(const float *x, float *out, long long rows, int C) {
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const float *src = x + row * C;
    float ss = 0.f;
    for (int c = lane; c < C; c += 32) { const float v = src[c]; ss += v * v; }
    ss = warp_sum(ss);
    const float den = fmaxf(sqrtf(ss), 1e-12f);           // F.normalize(p=2, eps=1e-12): v / max(|v|, eps)
    for (int c = lane; c < C; c += 32) out[row * C + c] = src[c] / den;
}

}  // namespace opb

extern "C" int op_linear_row_tiles(int rows_per_batch) { return opb::ceil_div(rows_per_batch, opb::LBM); }

extern "C" int op_linear(const float *A1, int K1, const float *A2, int K2, const float *W, const float *bias,
                         float *C, int B, int rows_per_batch, int Cout, int flags, int elu_cols,
                         const float *mean, const float *rstd, double *stats_partial, void *stream) {
    OPB_CHECK_ARG(A1 && W && C, "op_linear: null pointer");
    OPB_CHECK_ARG(K1 > 0 && K1 % opb::LBK == 0 && K2 >= 0 && K2 % opb::LBK == 0, "op_linear: K1=%d K2=%d must be multiples of 16", K1, K2);
    OPB_CHECK_ARG((K2 == 0) == (A2 == nullptr), "op_linear: A2/K2 mismatch");
    OPB_CHECK_ARG(Cout > 0 && Cout % opb::LBN == 0, "op_linear: Cout=%d must be a multiple of 128", Cout);
    OPB_CHECK_ARG(!(flags & OPB_LIN_NORM_RELU) || (mean && rstd), "op_linear: NORM_RELU needs mean/rstd");
    OPB_CHECK_ARG(!(flags & OPB_LIN_STATS) || stats_partial, "op_linear: STATS needs stats_partial");
    OPB_CHECK_ARG(B >= 0 && rows_per_batch >= 0 && B <= 65535, "op_linear: bad B=%d rows=%d", B, rows_per_batch);
    if (B == 0 || rows_per_batch == 0) return OPB_OK;
    opb::LinearParams p{A1, K1, A2, K2, W, bias, C, rows_per_batch, Cout, flags, elu_cols, mean, rstd, stats_partial,
                        opb::ceil_div(rows_per_batch, opb::LBM)};
    dim3 grid(p.row_tiles, Cout / opb::LBN, B);
    opb::k_linear_simt<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_instnorm_finalize(const double *stats_partial, int B, int row_tiles, int C, int rows_per_batch,
                                    float eps, float *mean, float *rstd, void *stream) {
    OPB_CHECK_ARG(stats_partial && mean && rstd, "op_instnorm_finalize: null pointer");
    OPB_CHECK_ARG(B > 0 && row_tiles > 0 && C > 0 && rows_per_batch > 0, "op_instnorm_finalize: bad shape");
    dim3 grid(opb::ceil_div(C, 128), B);
    opb::k_instnorm_finalize<<<grid, 128, 0, (cudaStream_t)stream>>>(stats_partial, row_tiles, C, rows_per_batch, eps, mean, rstd);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_l2_normalize_rows(const float *x, float *out, long long rows, int C, void *stream) {
    OPB_CHECK_ARG(x && out && C > 0 && rows >= 0, "op_l2_normalize_rows: bad argument");
    if (rows == 0) return OPB_OK;
    opb::k_l2_normalize_rows<<<opb::ceil_div(rows, 8), 256, 0, (cudaStream_t)stream>>>(x, out, rows, C);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
