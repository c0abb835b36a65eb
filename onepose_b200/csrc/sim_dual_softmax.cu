// (3) similarity + dual-softmax + mutual-NN -- reference GATs_SuperGlue.py:251-274 (fp32 CUDA-core version).
//
// conf[n,m] = softmax_n(s)[n,m] * softmax_m(s)[n,m],  s = <m2[n], m3[m]> / scale_factor.
// Both operands are unit vectors, so |s| <= 1/scale and a FIXED shift of 1/scale replaces the running
// max of an online softmax (no rescaling pass).  Pass 1 recomputes nothing twice: it produces only the
// row sums R_n and column sums C_m of e = exp(s - 1/scale).  Pass 2 recomputes the tile, writes
// conf = (e/R_n)*(e/C_m) to HBM exactly once and keeps per-row / per-column (max, argmax) as packed
// 64-bit keys (value bits high, inverted index low => ties resolve to the smallest index).  All
// cross-tile reductions go through per-tile partials folded in a fixed order (deterministic).
#include <cstring>
#include <stdlib.h>
#include "common.cuh"
#include "sim_tc.cuh"
#include "sim_tc16.cuh"

namespace opb {

constexpr int SBM = 128, SBN = 128, SBK = 16, SPAD = 4;

struct SimParams {
    const float *m2, *m3;
    int N2, N3, tiles_m, tiles_n;
    float inv_unused, scale, smax;
    float *rowsum_part, *colsum_part;        // [B][tiles_n][N2], [B][tiles_m][N3]
    const float *invR, *invC;                // [B][N2], [B][N3]
    unsigned long long *rowmax_part, *colmax_part;
    float *conf;
};

__device__ __forceinline__ unsigned long long pack_key(float v, int idx) {
    return ((unsigned long long)__float_as_uint(v) << 32) | (unsigned long long)(0xffffffffu - (unsigned)idx);
}

template <int PASS>
__global__ void __launch_bounds__(256)
k_sim_pass(const SimParams p) {
    __shared__ __align__(16) float As[SBK][SBM + SPAD];
    __shared__ __align__(16) float Ws[SBK][SBN + SPAD];
    __shared__ float cred[8][SBN];
    __shared__ unsigned long long kred[PASS == 2 ? 8 : 1][SBN];

    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int b = blockIdx.z, m0 = blockIdx.y * SBM, n0 = blockIdx.x * SBN;   // m0: query rows, n0: 3D columns
    const float *A = p.m2 + (long long)b * p.N2 * kD;
    const float *Bm = p.m3 + (long long)b * p.N3 * kD;
    const int lr = tid >> 2, lk = (tid & 3) * 4;
    float4 ra[2], rw[2];
    auto load_tile = [&](int k0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = m0 + lr + h * 64, c = n0 + lr + h * 64;
            ra[h] = r < p.N2 ? *reinterpret_cast<const float4 *>(A + (long long)r * kD + k0 + lk) : make_float4(0, 0, 0, 0);
            rw[h] = c < p.N3 ? *reinterpret_cast<const float4 *>(Bm + (long long)c * kD + k0 + lk) : make_float4(0, 0, 0, 0);
        }
    };
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    load_tile(0);
    for (int k0 = 0; k0 < kD; k0 += SBK) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = lr + h * 64;
            As[lk + 0][r] = ra[h].x; As[lk + 1][r] = ra[h].y; As[lk + 2][r] = ra[h].z; As[lk + 3][r] = ra[h].w;
            Ws[lk + 0][r] = rw[h].x; Ws[lk + 1][r] = rw[h].y; Ws[lk + 2][r] = rw[h].z; Ws[lk + 3][r] = rw[h].w;
        }
        __syncthreads();
        if (k0 + SBK < kD) load_tile(k0 + SBK);
#pragma unroll
        for (int k = 0; k < SBK; ++k) {
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

    // e = exp(s - smax), zero outside the matrix
    int rows[8], cols[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { rows[i] = m0 + (i >> 2) * 64 + ty * 4 + (i & 3); cols[i] = n0 + (i >> 2) * 64 + tx * 4 + (i & 3); }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j)
            acc[i][j] = (rows[i] < p.N2 && cols[j] < p.N3) ? expf(acc[i][j] / p.scale - p.smax) : 0.f;

    const int warp = tid >> 5;
    if (PASS == 1) {
        // row sums: fold 8 columns per thread, then the 16 tx lanes of a half-warp
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) s += acc[i][j];
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (tx == 0 && rows[i] < p.N2) p.rowsum_part[((long long)b * p.tiles_n + blockIdx.x) * p.N2 + rows[i]] = s;
        }
        // column sums: 8 rows per thread, pair of ty via shuffle, 8 warps via smem
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) s += acc[i][j];
            s += __shfl_xor_sync(0xffffffffu, s, 16);
            if ((tid & 16) == 0) cred[warp][(j >> 2) * 64 + tx * 4 + (j & 3)] = s;
        }
        __syncthreads();
        if (tid < SBN && n0 + tid < p.N3) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) s += cred[w][tid];
            p.colsum_part[((long long)b * p.tiles_m + blockIdx.y) * p.N3 + n0 + tid] = s;
        }
    } else {
        float ir[8], ic[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            ir[i] = rows[i] < p.N2 ? p.invR[(long long)b * p.N2 + rows[i]] : 0.f;
            ic[i] = cols[i] < p.N3 ? p.invC[(long long)b * p.N3 + cols[i]] : 0.f;
        }
        const bool vec = (p.N3 & 3) == 0;
        unsigned long long ckey[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) ckey[j] = 0ull;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            unsigned long long rkey = 0ull;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float c = (acc[i][j] * ic[j]) * (acc[i][j] * ir[i]);     // softmax(dim=1) * softmax(dim=2), :253
                acc[i][j] = c;
                if (rows[i] < p.N2 && cols[j] < p.N3) {
                    const unsigned long long kr = pack_key(c, cols[j]), kc = pack_key(c, rows[i]);
                    rkey = kr > rkey ? kr : rkey;
                    ckey[j] = kc > ckey[j] ? kc : ckey[j];
                }
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) { const unsigned long long ok = __shfl_xor_sync(0xffffffffu, rkey, o); rkey = ok > rkey ? ok : rkey; }
            if (rows[i] < p.N2) {
                if (tx == 0) p.rowmax_part[((long long)b * p.tiles_n + blockIdx.x) * p.N2 + rows[i]] = rkey;
                float *crow = p.conf + ((long long)b * p.N2 + rows[i]) * p.N3;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c0 = n0 + h * 64 + tx * 4;
                    if (vec && c0 + 3 < p.N3) {
                        *reinterpret_cast<float4 *>(crow + c0) = make_float4(acc[i][h * 4], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]);
                    } else {
#pragma unroll
                        for (int q = 0; q < 4; ++q) if (c0 + q < p.N3) crow[c0 + q] = acc[i][h * 4 + q];
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const unsigned long long ok = __shfl_xor_sync(0xffffffffu, ckey[j], 16);
            ckey[j] = ok > ckey[j] ? ok : ckey[j];
            if ((tid & 16) == 0) kred[warp][(j >> 2) * 64 + tx * 4 + (j & 3)] = ckey[j];
        }
        __syncthreads();
        if (tid < SBN && n0 + tid < p.N3) {
            unsigned long long k = 0ull;
#pragma unroll
            for (int w = 0; w < 8; ++w) k = kred[w][tid] > k ? kred[w][tid] : k;
            p.colmax_part[((long long)b * p.tiles_m + blockIdx.y) * p.N3 + n0 + tid] = k;
        }
    }
}

// inv[b][i] = 1 / sum_t part[b][t][i]   (as_log2: log2 of it, for the exponent-folded tcgen05 pass 2)
__global__ void k_sim_finalize_sums(const float *__restrict__ part, int tiles, int n, float *__restrict__ inv, int as_log2) {
    const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0.f;
    for (int t = 0; t < tiles; ++t) s += part[((long long)b * tiles + t) * n + i];
    // as_log2: a sum of zero (a frame of a ragged batch without valid rows) gives the neutral term, conf stays +0 instead of inf - inf
    inv[(long long)b * n + i] = as_log2 ? (s > 0.f ? -log2f(s) : 0.f) : 1.f / s;
}

__global__ void k_sim_finalize_max(const unsigned long long *__restrict__ part, int tiles, int n, int n_other,
                                   int *__restrict__ idx, float *__restrict__ val, const float *__restrict__ conf_rows,
                                   const int *__restrict__ n_counts = nullptr) {
    const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (n_counts && i >= n_counts[b]) {             // padding row of a ragged batch: its partials were never written
        idx[(long long)b * n + i] = 0; val[(long long)b * n + i] = 0.f;
        return;
    }
    unsigned long long k = 0ull;
    for (int t = 0; t < tiles; ++t) { const unsigned long long v = part[((long long)b * tiles + t) * n + i]; k = v > k ? v : k; }
    // an all-zero row / column of conf can select a padding slot of the last tile: every index is then a maximum, clamp it
    int id = (int)(0xffffffffu - (unsigned)(k & 0xffffffffull));
    id = id < n_other ? id : n_other - 1;
    if (conf_rows) {
        // tensor-core pass 2, rows: the key carries the exact maximum and the first column of the 16-column block that holds it;
        // the column is the first of those 16 scores equal to it (conf_rows = conf, row i of frame b)
        const float *r = conf_rows + ((long long)b * n + i) * n_other;
        const unsigned want = (unsigned)(k >> 32);
        int hit = id;
        for (int c = 15; c >= 0; --c) if (id + c < n_other && __float_as_uint(r[id + c]) == want) hit = id + c;
        id = hit;
    }
    idx[(long long)b * n + i] = id;
    val[(long long)b * n + i] = __uint_as_float((unsigned)(k >> 32));
}

// mutual nearest neighbour + threshold, GATs_SuperGlue.py:256-267
// The score of a pair is read back from conf itself (exact), the packed keys only select.
__global__ void k_mutual_select(const int *__restrict__ idx0, const float *__restrict__ conf, const int *__restrict__ idx1,
                                int N2, int N3, float thr, int64_t *__restrict__ matches0, int64_t *__restrict__ matches1,
                                float *__restrict__ scores0, float *__restrict__ scores1, const int *__restrict__ n2_counts = nullptr) {
    const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    const int *i0 = idx0 + (long long)b * N2, *i1 = idx1 + (long long)b * N3;
    const float *cb = conf + (long long)b * N2 * N3;
    const int n2 = n2_counts ? min(N2, n2_counts[b]) : N2;      // ragged batch: rows past the count are padding and never match
    if (i < N2) {
        const int j = i0[i];
        const bool mutual = i < n2 && i1[j] == i;
        const float ms = mutual ? cb[(long long)i * N3 + j] : 0.f;
        scores0[(long long)b * N2 + i] = ms;
        matches0[(long long)b * N2 + i] = (mutual && ms > thr) ? (int64_t)j : (int64_t)-1;
    }
    if (i < N3) {
        const int n = i1[i];                       // best query for this 3D point
        const bool mutual1 = n < n2 && i0[n] == i;
        const bool mutual0_n = i1[i0[n]] == n;
        const float ms0_n = mutual0_n ? cb[(long long)n * N3 + i0[n]] : 0.f;
        scores1[(long long)b * N3 + i] = mutual1 ? ms0_n : 0.f;
        const bool valid0_n = mutual0_n && ms0_n > thr;
        matches1[(long long)b * N3 + i] = (mutual1 && valid0_n) ? (int64_t)n : (int64_t)-1;
    }
}

struct SimWorkspace {
    float *rowsum_part, *colsum_part, *invR, *invC, *val0, *val1;
    float *colsum4;              // sixteen-warp form: column sums per (row tile, TMEM lane quadrant)  [B][tm * 4][N3]
    uint32_t *colkey4;           // sixteen-warp form: column keys per (row tile, TMEM lane quadrant)  [B][tm * 4][N3]
    __half *m2h, *m2l, *m3h, *m3l;
    unsigned long long *rowmax_part, *colmax_part;
    int *idx0, *idx1;
    size_t bytes;
};

static SimWorkspace carve(void *base, int B, int N2, int N3) {
    const size_t tm = ceil_div(N2, SBM);
    // partials per query row: two per 128-column tile (eight-warp forms) or four per 256-column tile (sixteen-warp form)
    const size_t tn = ceil_div(N3, SBN) > 2 * (size_t)ceil_div(N3, 256) ? ceil_div(N3, SBN) : 2 * (size_t)ceil_div(N3, 256);
    SimWorkspace w;
    size_t off = 0;
    auto take = [&](size_t bytes) { void *p = base ? (char *)base + off : nullptr; off += align_up(bytes, 256); return p; };
    w.rowmax_part = (unsigned long long *)take((size_t)B * tn * 2 * N2 * 8);   // x2: the tcgen05 epilogue splits a tile's columns over two warps
    w.colmax_part = (unsigned long long *)take((size_t)B * tm * N3 * 8);
    w.rowsum_part = (float *)take((size_t)B * tn * 2 * N2 * 4);
    w.colsum_part = (float *)take((size_t)B * tm * N3 * 4);
    w.colsum4 = (float *)take((size_t)B * tm * 4 * N3 * 4);
    w.colkey4 = (uint32_t *)take((size_t)B * tm * 4 * N3 * 4);
    w.invR = (float *)take((size_t)B * N2 * 4);
    w.invC = (float *)take((size_t)B * N3 * 4);
    w.val0 = (float *)take((size_t)B * N2 * 4);
    w.val1 = (float *)take((size_t)B * N3 * 4);
    w.idx0 = (int *)take((size_t)B * N2 * 4);
    w.idx1 = (int *)take((size_t)B * N3 * 4);
    w.m2h = (__half *)take((size_t)B * N2 * kD * 2);
    w.m2l = (__half *)take((size_t)B * N2 * kD * 2);
    w.m3h = (__half *)take((size_t)B * N3 * kD * 2);
    w.m3l = (__half *)take((size_t)B * N3 * kD * 2);
    w.bytes = off;
    return w;
}

}  // namespace opb

extern "C" size_t op_sim_dual_softmax_workspace_bytes(int B, int N2, int N3) {
    return opb::carve(nullptr, B > 0 ? B : 1, N2 > 0 ? N2 : 1, N3 > 0 ? N3 : 1).bytes;
}

// planes (optional): {m2 hi fp16, m2 bytes, m3 hi fp16, m3 bytes} -- operand planes some producer has already written (the final projection
// with OPB_LIN_UNIT_PLANES8); bytes = lo e5m2 plane followed by the hi e4m3 plane.  m2 / m3 are then unused.
static int sim_impl(bool tensor_cores, bool normalize, const float *m2, const float *m3, int B, int N2, int N3, float scale_factor,
                    float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                    float *scores0, float *scores1, void *workspace, size_t workspace_bytes,
                    void *stream, const void *const *planes = nullptr, const int *n2_counts = nullptr) {
    using namespace opb;
    OPB_CHECK_ARG((planes || (m2 && m3)) && conf && matches0 && matches1 && scores0 && scores1 && workspace, "op_sim_dual_softmax: null pointer");
    OPB_CHECK_ARG(B > 0 && N2 > 0 && N3 > 0 && B <= 65535, "op_sim_dual_softmax: bad shape B=%d N2=%d N3=%d", B, N2, N3);
    OPB_CHECK_ARG(scale_factor > 0.f, "op_sim_dual_softmax: scale_factor must be positive");
    SimWorkspace w = carve(workspace, B, N2, N3);
    if (workspace_bytes < w.bytes) { set_error("op_sim_dual_softmax: workspace %zu < %zu bytes", workspace_bytes, w.bytes); return OPB_ERR_WORKSPACE; }
    cudaStream_t st = (cudaStream_t)stream;
    SimParams p{};
    p.m2 = m2; p.m3 = m3; p.N2 = N2; p.N3 = N3; p.tiles_m = ceil_div(N2, SBM); p.tiles_n = ceil_div(N3, SBN);
    p.scale = scale_factor; p.smax = 1.f / scale_factor;
    p.rowsum_part = w.rowsum_part; p.colsum_part = w.colsum_part; p.invR = w.invR; p.invC = w.invC;
    p.rowmax_part = w.rowmax_part; p.colmax_part = w.colmax_part; p.conf = conf;
    OPB_CHECK_ARG(p.tiles_m <= 65535, "op_sim_dual_softmax: N2 too large");
    int row_parts = p.tiles_n;                       // partials per query row (column tiles or column splits)
    if (!tensor_cores) {
        dim3 grid(p.tiles_n, p.tiles_m, B);
        k_sim_pass<1><<<grid, 256, 0, st>>>(p);
        OPB_CHECK_LAUNCH();
        k_sim_finalize_sums<<<dim3(ceil_div(N2, 256), B), 256, 0, st>>>(w.rowsum_part, row_parts, N2, w.invR, 0);
        k_sim_finalize_sums<<<dim3(ceil_div(N3, 256), B), 256, 0, st>>>(w.colsum_part, p.tiles_m, N3, w.invC, 0);
        OPB_CHECK_LAUNCH();
        k_sim_pass<2><<<grid, 256, 0, st>>>(p);
        OPB_CHECK_LAUNCH();
    } else {
        // split fp32 unit rows into scaled fp16 hi/lo planes, describe them to TMA, run the two tcgen05 passes
        const long long n2e = (long long)B * N2 * kD, n3e = (long long)B * N3 * kD;
        // One CTA per row tile pair by default (cta_group::2, M256 x N256 tiles): each CTA keeps its 128 query rows resident and stages half
        // of every B k-block, which halves the L2 -> SM stream that bounds the one-CTA form (OPB_SIM_CG=1, M128 x N128).  In the pair form
        // the two cross terms of the split-precision product run as fp8 MMAs (OPB_SIM_MMA=f16 keeps three fp16 MMAs per product).
        static const int cg = [] { const char *e = getenv("OPB_SIM_CG"); return (e && e[0] == '1') ? 1 : 2; }();
        static const int f8 = [] { const char *e = getenv("OPB_SIM_MMA"); return (e && !strcmp(e, "f16")) ? 0 : 1; }() && cg == 2;
        // fp8 mode: the 2-byte lo area holds the e5m2 lo plane (first half) and the e4m3 hi plane (second half)
        uint8_t *m2x = (uint8_t *)w.m2l + (size_t)B * N2 * kD, *m3x = (uint8_t *)w.m3l + (size_t)B * N3 * kD;
        const long long r2 = (long long)B * N2, r3 = (long long)B * N3;
        if (planes) {
            if (!f8) { set_error("op_sim_dual_softmax_planes: the byte planes belong to the pair form with fp8 cross terms (unset OPB_SIM_CG / OPB_SIM_MMA)"); return OPB_ERR_ARG; }
            w.m2h = (__half *)planes[0]; w.m2l = (__half *)planes[1]; m2x = (uint8_t *)planes[1] + (size_t)B * N2 * kD;
            w.m3h = (__half *)planes[2]; w.m3l = (__half *)planes[3]; m3x = (uint8_t *)planes[3] + (size_t)B * N3 * kD;
        } else if (normalize && f8) {
            simtc::k_normalize_split_planes<1><<<ceil_div(r2, 8), 256, 0, st>>>(m2, w.m2h, w.m2l, m2x, r2, simtc::kOperandScale);
            simtc::k_normalize_split_planes<1><<<ceil_div(r3, 8), 256, 0, st>>>(m3, w.m3h, w.m3l, m3x, r3, simtc::kOperandScale);
        } else if (normalize) {
            simtc::k_normalize_split_planes<0><<<ceil_div(r2, 8), 256, 0, st>>>(m2, w.m2h, w.m2l, m2x, r2, simtc::kOperandScale);
            simtc::k_normalize_split_planes<0><<<ceil_div(r3, 8), 256, 0, st>>>(m3, w.m3h, w.m3l, m3x, r3, simtc::kOperandScale);
        } else if (f8) {
            simtc::k_split_planes<1><<<ceil_div(n2e / 4, 256), 256, 0, st>>>(m2, w.m2h, w.m2l, m2x, n2e, simtc::kOperandScale);
            simtc::k_split_planes<1><<<ceil_div(n3e / 4, 256), 256, 0, st>>>(m3, w.m3h, w.m3l, m3x, n3e, simtc::kOperandScale);
        } else {
            simtc::k_split_planes<0><<<ceil_div(n2e / 4, 256), 256, 0, st>>>(m2, w.m2h, w.m2l, m2x, n2e, simtc::kOperandScale);
            simtc::k_split_planes<0><<<ceil_div(n3e / 4, 256), 256, 0, st>>>(m3, w.m3h, w.m3l, m3x, n3e, simtc::kOperandScale);
        }
        OPB_CHECK_LAUNCH();
        CUtensorMap t2h, t2l, t2x, t3h, t3l, t3x, tconf;
        int rc;
        if ((rc = make_tmap_f16_3d(&t2h, w.m2h, B, N2, kD, simtc::BM)) || (rc = make_tmap_f16_3d(&t3h, w.m3h, B, N3, kD, 128))) return rc;
        if (f8) {
            if ((rc = make_tmap_u8_3d(&t2l, w.m2l, B, N2, kD, simtc::BM)) || (rc = make_tmap_u8_3d(&t2x, m2x, B, N2, kD, simtc::BM)) ||
                (rc = make_tmap_u8_3d(&t3l, w.m3l, B, N3, kD, 128)) || (rc = make_tmap_u8_3d(&t3x, m3x, B, N3, kD, 128)))
                return rc;
        } else {
            if ((rc = make_tmap_f16_3d(&t2l, w.m2l, B, N2, kD, simtc::BM)) || (rc = make_tmap_f16_3d(&t3l, w.m3l, B, N3, kD, 128))) return rc;
            t2x = t2l; t3x = t3l;
        }
        // sixteen epilogue warps (sim_tc16.cuh) are the production form of the pair kernels; OPB_SIM_EW=8 selects the eight-warp form for A/B runs
        static const bool ew16 = [] { const char *e = getenv("OPB_SIM_EW"); return !(e && e[0] == '8'); }();
        if (cg == 2 && ew16) {
            const int tiles_n = ceil_div(N3, simtc16::BN);
            int splits = ceil_div(2 * 148, (long long)B * p.tiles_m);
            splits = splits < 1 ? 1 : (splits > tiles_n ? tiles_n : splits);
            simtc16::Params q{};
            q.N2 = N2; q.N3 = N3; q.tiles_n = tiles_n; q.tiles_m = p.tiles_m; q.tiles_per_split = ceil_div(tiles_n, splits);
            splits = ceil_div(tiles_n, q.tiles_per_split);
            const double log2e = 1.4426950408889634;
            q.c1 = (float)(log2e / ((double)scale_factor * simtc::kOperandScale * simtc::kOperandScale));
            q.c0 = (float)(log2e / (double)scale_factor);
            q.c2 = 2.f * q.c1;
            q.tma_store = (N3 % 4 == 0) && ((uintptr_t)conf % 16 == 0);
            if (q.tma_store) { if ((rc = make_tmap_f32_3d_store(&tconf, conf, B, N2, N3, 32))) return rc; } else tconf = t2h;
            { const char *e = getenv("OPB_SIM_DBG"); q.dbg = e ? atoi(e) : 0; }
            q.rowsum_part = w.rowsum_part; q.colsum_part = w.colsum4; q.lgR = w.invR; q.lgC = w.invC;
            q.rowmax_part = w.rowmax_part; q.colkey_part = w.colkey4; q.conf = conf;
            q.n2_counts = n2_counts;
            static bool configured16 = false;
            if (!configured16) {
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<1>()));
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<2>()));
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<1>()));
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<2>()));
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<2, 1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<2>()));
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<2, 1, 1, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<2, 2, 0>()));
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<2, 1, 1, 3, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<2, 3, 0>()));
                OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc16::k_sim16<2, 1, 1, 2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc16::smem_bytes<2, 2, 1>()));
                configured16 = true;
            }
            cudaLaunchConfig_t cfg = {};
            cudaLaunchAttribute attr[1];
            cfg.gridDim = dim3(ceil_div(p.tiles_m, 2) * 2, splits, B);        // a ghost row tile completes the last pair
            cfg.blockDim = dim3(simtc16::THREADS);
            cfg.stream = st;
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            cfg.dynamicSmemBytes = simtc16::smem_bytes<1>();
            if (f8) OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<1, 1>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
            else OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<1, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
            k_sim_finalize_sums<<<dim3(ceil_div(N2, 256), B), 256, 0, st>>>(w.rowsum_part, tiles_n * 4, N2, w.invR, 1);
            k_sim_finalize_sums<<<dim3(ceil_div(N3, 256), B), 256, 0, st>>>(w.colsum4, p.tiles_m * 4, N3, w.invC, 1);
            OPB_CHECK_LAUNCH();
            cfg.dynamicSmemBytes = simtc16::smem_bytes<2>();
            // conf leaves as TMA stores of 16 rows x 128 bytes (OPB_SIM_STORE=narrow: 32 rows x 64 bytes, the eight-warp form's boxes)
            static const bool wide = [] { const char *e = getenv("OPB_SIM_STORE"); return !(e && e[0] == 'n'); }();
            CUtensorMap tconf_w = tconf;
            if (f8 && wide && q.tma_store) { if ((rc = make_tmap_f32_3d_store_wide(&tconf_w, conf, B, N2, N3, 16))) return rc; }
            // pass-2 shared memory plan (OPB_SIM_P2): "4s" = four ring units + one staging buffer per warp, "2d" = two units + two buffers,
            // "2s" / "3s" = fewer units with one buffer (A/B runs of the ring depth alone)
            static const int p2plan = [] { const char *e = getenv("OPB_SIM_P2"); return !e ? 0 : (!strcmp(e, "2d") ? 1 : !strcmp(e, "2s") ? 2 : !strcmp(e, "3s") ? 3 : 0); }();
            if (f8 && wide && p2plan == 1) { cfg.dynamicSmemBytes = simtc16::smem_bytes<2, 2, 1>(); OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<2, 1, 1, 2, 1>, t2h, t2l, t2x, t3h, t3l, t3x, tconf_w, q)); }
            else if (f8 && wide && p2plan == 2) { cfg.dynamicSmemBytes = simtc16::smem_bytes<2, 2, 0>(); OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<2, 1, 1, 2, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf_w, q)); }
            else if (f8 && wide && p2plan == 3) { cfg.dynamicSmemBytes = simtc16::smem_bytes<2, 3, 0>(); OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<2, 1, 1, 3, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf_w, q)); }
            else if (f8 && wide) OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<2, 1, 1>, t2h, t2l, t2x, t3h, t3l, t3x, tconf_w, q));
            else if (f8) OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<2, 1>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
            else OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc16::k_sim16<2, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
            k_sim_finalize_max<<<dim3(ceil_div(N2, 256), B), 256, 0, st>>>(w.rowmax_part, splits * 4, N2, N3, w.idx0, w.val0, conf, n2_counts);
            simtc16::k_sim_finalize_colkeys<<<dim3(ceil_div(N3, 256), B), 256, 0, st>>>(w.colkey4, p.tiles_m * 4, N3, N2, w.idx1, w.val1);
            OPB_CHECK_LAUNCH();
            const int nmax16 = N2 > N3 ? N2 : N3;
            k_mutual_select<<<dim3(ceil_div(nmax16, 256), B), 256, 0, st>>>(w.idx0, conf, w.idx1, N2, N3, match_threshold, matches0, matches1, scores0, scores1, n2_counts);
            OPB_CHECK_LAUNCH();
            return OPB_OK;
        }
        if (n2_counts) { set_error("op_sim_dual_softmax_planes_ragged: per-frame row counts exist in the sixteen-warp pair form only (unset OPB_SIM_CG / OPB_SIM_EW)"); return OPB_ERR_ARG; }
        const int tiles_n = ceil_div(N3, 128 * cg);          // column tiles of the tensor-core passes: 128 wide, 256 for a CTA pair
        // enough CTAs for ~2 waves over 148 SMs: split the column tiles when frames x row tiles is small
        int splits = ceil_div(2 * 148, (long long)B * p.tiles_m);
        splits = splits < 1 ? 1 : (splits > tiles_n ? tiles_n : splits);
        simtc::Params q{};
        q.N2 = N2; q.N3 = N3; q.tiles_n = tiles_n; q.tiles_m = p.tiles_m; q.tiles_per_split = ceil_div(tiles_n, splits);
        splits = ceil_div(tiles_n, q.tiles_per_split);
        row_parts = splits * 2;
        const double log2e = 1.4426950408889634;
        q.c1 = (float)(log2e / ((double)scale_factor * simtc::kOperandScale * simtc::kOperandScale));
        q.c0 = (float)(log2e / (double)scale_factor);
        q.c2 = 2.f * q.c1;
        q.tma_store = (N3 % 4 == 0) && ((uintptr_t)conf % 16 == 0);
        if (q.tma_store) { if ((rc = make_tmap_f32_3d_store(&tconf, conf, B, N2, N3, 32))) return rc; } else tconf = t2h;
        { const char *e = getenv("OPB_SIM_DBG"); q.dbg = e ? atoi(e) : 0; }
        q.rowsum_part = w.rowsum_part; q.colsum_part = w.colsum_part; q.lgR = w.invR; q.lgC = w.invC;
        q.rowmax_part = w.rowmax_part; q.colmax_part = w.colmax_part; q.conf = conf;
        static bool configured = false;
        if (!configured) {
            OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc::k_sim_tc<1, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc::SMEM_BYTES));
            OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc::k_sim_tc<2, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc::SMEM_BYTES));
            OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc::k_sim_tc<1, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc::SMEM_BYTES));
            OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc::k_sim_tc<2, 2, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc::SMEM_BYTES));
            OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc::k_sim_tc<1, 2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc::SMEM_BYTES));
            OPB_CHECK_CUDA(cudaFuncSetAttribute(simtc::k_sim_tc<2, 2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, simtc::SMEM_BYTES));
            configured = true;
        }
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute attr[1];
        cfg.gridDim = dim3(ceil_div(p.tiles_m, cg) * cg, splits, B);       // a ghost row tile completes the last pair
        cfg.blockDim = dim3(320);
        cfg.dynamicSmemBytes = simtc::SMEM_BYTES;
        cfg.stream = st;
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = cg; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        if (f8) OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc::k_sim_tc<1, 2, 1>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
        else if (cg == 2) OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc::k_sim_tc<1, 2, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
        else OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc::k_sim_tc<1, 1, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
        k_sim_finalize_sums<<<dim3(ceil_div(N2, 256), B), 256, 0, st>>>(w.rowsum_part, tiles_n * 2, N2, w.invR, 1);
        k_sim_finalize_sums<<<dim3(ceil_div(N3, 256), B), 256, 0, st>>>(w.colsum_part, p.tiles_m, N3, w.invC, 1);
        OPB_CHECK_LAUNCH();
        if (f8) OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc::k_sim_tc<2, 2, 1>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
        else if (cg == 2) OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc::k_sim_tc<2, 2, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
        else OPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, simtc::k_sim_tc<2, 1, 0>, t2h, t2l, t2x, t3h, t3l, t3x, tconf, q));
    }
    k_sim_finalize_max<<<dim3(ceil_div(N2, 256), B), 256, 0, st>>>(w.rowmax_part, row_parts, N2, N3, w.idx0, w.val0, tensor_cores ? conf : nullptr);
    k_sim_finalize_max<<<dim3(ceil_div(N3, 256), B), 256, 0, st>>>(w.colmax_part, p.tiles_m, N3, N2, w.idx1, w.val1, nullptr);
    OPB_CHECK_LAUNCH();
    const int nmax = N2 > N3 ? N2 : N3;
    k_mutual_select<<<dim3(ceil_div(nmax, 256), B), 256, 0, st>>>(w.idx0, conf, w.idx1, N2, N3, match_threshold,
                                                                  matches0, matches1, scores0, scores1);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_sim_dual_softmax(const float *m2, const float *m3, int B, int N2, int N3, float scale_factor,
                                   float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                                   float *scores0, float *scores1, void *workspace, size_t workspace_bytes, void *stream) {
    return sim_impl(true, false, m2, m3, B, N2, N3, scale_factor, match_threshold, conf, matches0, matches1, scores0, scores1,
                    workspace, workspace_bytes, stream);
}

extern "C" int op_sim_dual_softmax_simt(const float *m2, const float *m3, int B, int N2, int N3, float scale_factor,
                                        float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                                        float *scores0, float *scores1, void *workspace, size_t workspace_bytes, void *stream) {
    return sim_impl(false, false, m2, m3, B, N2, N3, scale_factor, match_threshold, conf, matches0, matches1, scores0, scores1,
                    workspace, workspace_bytes, stream);
}

extern "C" int op_sim_dual_softmax_rows(const float *x2, const float *x3, int B, int N2, int N3, float scale_factor,
                                        float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                                        float *scores0, float *scores1, void *workspace, size_t workspace_bytes, void *stream) {
    return sim_impl(true, true, x2, x3, B, N2, N3, scale_factor, match_threshold, conf, matches0, matches1, scores0, scores1,
                    workspace, workspace_bytes, stream);
}

extern "C" int op_sim_dual_softmax_planes_ragged(const void *m2_hi, const void *m2_bytes, const void *m3_hi, const void *m3_bytes, int B, int N2, int N3,
                                                 const int *n2_counts, float scale_factor, float match_threshold, float *conf, int64_t *matches0,
                                                 int64_t *matches1, float *scores0, float *scores1, void *workspace, size_t workspace_bytes, void *stream) {
    OPB_CHECK_ARG(m2_hi && m2_bytes && m3_hi && m3_bytes, "op_sim_dual_softmax_planes: null plane");
    OPB_CHECK_ARG((((uintptr_t)m2_hi | (uintptr_t)m2_bytes | (uintptr_t)m3_hi | (uintptr_t)m3_bytes) & 15) == 0, "op_sim_dual_softmax_planes: planes must be 16-byte aligned");
    const void *planes[4] = {m2_hi, m2_bytes, m3_hi, m3_bytes};
    return sim_impl(true, false, nullptr, nullptr, B, N2, N3, scale_factor, match_threshold, conf, matches0, matches1, scores0, scores1,
                    workspace, workspace_bytes, stream, planes, n2_counts);
}

extern "C" int op_sim_dual_softmax_planes(const void *m2_hi, const void *m2_bytes, const void *m3_hi, const void *m3_bytes, int B, int N2, int N3,
                                          float scale_factor, float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                                          float *scores0, float *scores1, void *workspace, size_t workspace_bytes, void *stream) {
    return op_sim_dual_softmax_planes_ragged(m2_hi, m2_bytes, m3_hi, m3_bytes, B, N2, N3, nullptr, scale_factor, match_threshold, conf, matches0, matches1,
                                             scores0, scores1, workspace, workspace_bytes, stream);
}
