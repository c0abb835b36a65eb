// SURVEY.md 8(f) row f3 -- SuperPoint's convolution stack (superpoint.py:147-166,177-206,231-234) as tcgen05 implicit GEMMs at
// fp32-class accuracy, on the same split-precision fp16 planes as the matcher's linear layers (tc_common.cuh, linear_tc.cuh).
//
//  k_conv_tc<BN>     : 3x3 (pad 1) or 1x1 convolution, NHWC.  out[pixel, co] = sum_{tap, ci} in[pixel + tap, ci] * w[tap, co, ci] is a GEMM
//                      whose A tile for (tap, 64-channel block) is the activation tile shifted by the tap: one 4-D TMA box
//                      {64 channels, TW pixels, TH rows} per stage, the image border zero-filled by the TMA unit.  Three MMAs per
//                      product (Ah*Wh + Al*Wh + Ah*Wl) accumulate in fp32 TMEM; the epilogue adds the bias, applies ReLU, and either
//                      re-splits into the planes the next layer reads or stores fp32 (the two heads' last layers).
//  k_conv1a_planes   : the first layer (one input channel, K = 9) on the CUDA cores, fp32, straight from the image into planes.
//  k_pool2_planes    : 2x2 max-pool on planes (max over hi + lo; the winning pair is copied).
#include <cstdlib>

#include "common.cuh"
#include "tc_common.cuh"

namespace opb {
namespace convtc {

constexpr int BM = 128, BK = 64;
constexpr int A_PLANE = BM * BK * 2;                  // 16384
constexpr int EPI_WARPS = 8, THREADS = (EPI_WARPS + 2) * 32;
constexpr int STG_BUF = 32 * 64;
constexpr int STG_BYTES = EPI_WARPS * 2 * STG_BUF;    // 32768
constexpr int BIAS_FLOATS = 256;
constexpr int MAX_STAGES = 4;
constexpr float kScale = 64.f;                        // planes hold 64 * value (activations and weights)

template <int BN>
struct Cfg {
    static constexpr int STAGES = BN == 64 ? 4 : 3;
    static constexpr int W_PLANE = BN * BK * 2;
    static constexpr int HALF_BYTES = A_PLANE + W_PLANE;
    static constexpr int STAGE_BYTES = 2 * HALF_BYTES;                 // 49152 / 65536
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + STG_BYTES + BIAS_FLOATS * 4 + 256;   // 230656
    static constexpr int TMEM_COLS = 2 * BN;
};

struct Params {
    int B, H, W, Cin, Cout, ntaps, relu;
    int TW, TH, tiles_x, tiles_y, col_tiles, total_tiles;
    int y32_stride;
    const float *bias;
    int has_planes, has_f32;
    int dbg;                       // experiment switches (env OPB_CONV_DBG; timing only, results are garbage): 1 = no loads, 2 = only the hi*hi MMA, 4 = no stores, 8 = no epilogue arithmetic or stores
};

struct Barriers {
    uint64_t full_hi[MAX_STAGES], full_lo[MAX_STAGES], empty[MAX_STAGES], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

// 32 fp32 values already scaled by 64 -> (hi, lo) fp16 pairs; saturated so that an activation beyond fp16's range (> 1023) cannot
// turn into inf
__device__ __forceinline__ void split_planes32(const float *vs, uint32_t *h, uint32_t *l) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float a = fminf(fmaxf(vs[2 * i], -65000.f), 65000.f), b = fminf(fmaxf(vs[2 * i + 1], -65000.f), 65000.f);
        const __half2 hi = __floats2half2_rn(a, b);
        const float2 hf = __half22float2(hi);
        const __half2 lo = __floats2half2_rn(a - hf.x, b - hf.y);
        h[i] = *reinterpret_cast<const uint32_t *>(&hi);
        l[i] = *reinterpret_cast<const uint32_t *>(&lo);
    }
}

template <int BN>
__global__ void __launch_bounds__(THREADS, 1)
k_conv_tc(const __grid_constant__ CUtensorMap tm_xh, const __grid_constant__ CUtensorMap tm_xl, const __grid_constant__ CUtensorMap tm_wh,
          const __grid_constant__ CUtensorMap tm_wl, const __grid_constant__ CUtensorMap tm_yh, const __grid_constant__ CUtensorMap tm_yl,
          const __grid_constant__ CUtensorMap tm_y32, const Params p) {
    constexpr int STAGES = Cfg<BN>::STAGES, HALF_BYTES = Cfg<BN>::HALF_BYTES, STAGE_BYTES = Cfg<BN>::STAGE_BYTES;
    extern __shared__ __align__(1024) uint8_t smem[];
    if (tc::smem_u32(smem) & 1023u) __trap();
    uint8_t *s_stg = smem + STAGES * STAGE_BYTES;
    float *s_bias = reinterpret_cast<float *>(smem + STAGES * STAGE_BYTES + STG_BYTES);
    Barriers *bar = reinterpret_cast<Barriers *>(smem + STAGES * STAGE_BYTES + STG_BYTES + BIAS_FLOATS * 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkb = p.Cin / BK;

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; ++i) { tc::mbar_init(&bar->full_hi[i], 1); tc::mbar_init(&bar->full_lo[i], 1); tc::mbar_init(&bar->empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&bar->acc_full[i], 1); tc::mbar_init(&bar->acc_empty[i], EPI_WARPS); }
        tc::fence_barrier_init();
    }
    for (int i = threadIdx.x; i < BIAS_FLOATS; i += THREADS) s_bias[i] = (p.bias && i < p.Cout) ? kScale * p.bias[i] : 0.f;
    if (warp == 9) tc::tmem_alloc<Cfg<BN>::TMEM_COLS>(&bar->tmem_base);
    if (warp == 8 && lane == 0) {
        tc::tma_prefetch_desc(&tm_xh); tc::tma_prefetch_desc(&tm_xl); tc::tma_prefetch_desc(&tm_wh); tc::tma_prefetch_desc(&tm_wl);
        if (p.has_planes) { tc::tma_prefetch_desc(&tm_yh); tc::tma_prefetch_desc(&tm_yl); }
        if (p.has_f32) tc::tma_prefetch_desc(&tm_y32);
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = bar->tmem_base;

    // tile t -> (image b, tile row, tile column, output-channel tile); the channel tile is the fastest index
    auto decode = [&](int t, int &b, int &y0, int &x0, int &ct) {
        ct = t % p.col_tiles; int r = t / p.col_tiles;
        x0 = (r % p.tiles_x) * p.TW; r /= p.tiles_x;
        y0 = (r % p.tiles_y) * p.TH; b = r / p.tiles_y;
    };

    if (warp == 8) {
        // ================= TMA producer =================
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
                int b, y0, x0, ct; decode(t, b, y0, x0, ct);
                for (int tap = 0; tap < p.ntaps; ++tap) {
                    const int dy = p.ntaps == 9 ? tap / 3 - 1 : 0, dx = p.ntaps == 9 ? tap % 3 - 1 : 0;
                    for (int kb = 0; kb < nkb; ++kb, ++it) {
                        const uint32_t stage = it % STAGES, ph = (it / STAGES) & 1;
                        uint8_t *st = smem + stage * STAGE_BYTES;
                        tc::mbar_wait(&bar->empty[stage], ph ^ 1);
                        tc::mbar_arrive_expect_tx(&bar->full_hi[stage], HALF_BYTES);
                        tc::tma_load_4d(st, &tm_xh, &bar->full_hi[stage], kb * BK, x0 + dx, y0 + dy, b);
                        tc::tma_load_3d(st + A_PLANE, &tm_wh, &bar->full_hi[stage], kb * BK, ct * BN, tap);
                        tc::mbar_arrive_expect_tx(&bar->full_lo[stage], HALF_BYTES);
                        tc::tma_load_4d(st + HALF_BYTES, &tm_xl, &bar->full_lo[stage], kb * BK, x0 + dx, y0 + dy, b);
                        tc::tma_load_3d(st + HALF_BYTES + A_PLANE, &tm_wl, &bar->full_lo[stage], kb * BK, ct * BN, tap);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = tc::umma_idesc_f16(BM, BN);
            const uint32_t base = tc::smem_u32(smem);
            const int nk = p.ntaps * nkb;
            uint32_t it = 0;
            int j = 0;
            for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++j) {
                const uint32_t as = j & 1, aph = (j >> 1) & 1;
                tc::mbar_wait(&bar->acc_empty[as], aph ^ 1);
                tc::tcgen05_fence_after();
                const uint32_t d = tmem_base + as * BN;
                for (int kk = 0; kk < nk; ++kk, ++it) {
                    const uint32_t stage = it % STAGES, ph = (it / STAGES) & 1;
                    const uint32_t sb = base + stage * STAGE_BYTES;
                    const uint64_t ah = tc::umma_desc_k128(sb), wh = tc::umma_desc_k128(sb + A_PLANE);
                    const uint64_t al = tc::umma_desc_k128(sb + HALF_BYTES), wl = tc::umma_desc_k128(sb + HALF_BYTES + A_PLANE);
                    tc::mbar_wait(&bar->full_hi[stage], ph);
                    tc::tcgen05_fence_after();
#pragma unroll
                    for (int k = 0; k < 4; ++k) tc::umma_f16(d, ah + 2 * k, wh + 2 * k, idesc, (kk | k) != 0);
                    tc::mbar_wait(&bar->full_lo[stage], ph);
                    tc::tcgen05_fence_after();
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        tc::umma_f16(d, al + 2 * k, wh + 2 * k, idesc, 1);
                        tc::umma_f16(d, ah + 2 * k, wl + 2 * k, idesc, 1);
                    }
                    tc::umma_commit(&bar->empty[stage]);
                }
                tc::umma_commit(&bar->acc_full[as]);
            }
        }
        __syncwarp();
    } else {
        // ================= epilogue (warps 0-7): warp w owns TMEM lanes 32*(w%4).., i.e. 32 consecutive pixels of one tile row =================
        const int quad = warp & 3, half = warp >> 2;
        const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
        uint32_t nst = 0;
        auto stage_store = [&](const CUtensorMap *tm, int col, const uint4 *data, int plane_row, int x) {
            uint8_t *buf = s_stg + (warp * 2 + (nst & 1)) * STG_BUF;
            if (lane == 0) tc::tma_store_wait_read1();
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4 *>(buf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4)) = data[q];
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) { tc::tma_store_3d(tm, buf, col, x, plane_row); tc::tma_store_commit(); }
            ++nst;
        };
        constexpr int CHUNKS = BN / 64;                           // 32-column chunks per warp
        int j = 0;
        for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++j) {
            int b, y0, x0, ct; decode(t, b, y0, x0, ct);
            const uint32_t as = j & 1, aph = (j >> 1) & 1;
            const int m0 = quad * 32;
            const int y = y0 + m0 / p.TW, xs = x0 + m0 % p.TW;    // this warp's 32 pixels: row y, columns xs .. xs+31
            const bool y_ok = y < p.H;
            tc::mbar_wait(&bar->acc_full[as], aph);
            tc::tcgen05_fence_after();
#pragma unroll 1
            for (int ch = 0; ch < CHUNKS; ++ch) {
                const int ccol = half * (BN / 2) + ch * 32;
                const int c0 = ct * BN + ccol;
                float v[32];
                tc::tmem_ld_32x32(lane_base + as * BN + ccol, v);
                tc::tmem_ld_wait();
                if (ch == CHUNKS - 1) {
                    tc::tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(&bar->acc_empty[as]);
                }
                if (!y_ok || xs >= p.W || c0 >= p.Cout) continue;
                // acc holds 4096 * value; work on vs = 64 * value, the scale of the planes
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float4 bq = *reinterpret_cast<const float4 *>(s_bias + c0 + 4 * q);
                    v[4 * q] = fmaf(v[4 * q], 1.f / kScale, bq.x); v[4 * q + 1] = fmaf(v[4 * q + 1], 1.f / kScale, bq.y);
                    v[4 * q + 2] = fmaf(v[4 * q + 2], 1.f / kScale, bq.z); v[4 * q + 3] = fmaf(v[4 * q + 3], 1.f / kScale, bq.w);
                }
                if (p.relu) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
                }
                if (p.has_planes) {
                    uint32_t h[16], l[16];
                    split_planes32(v, h, l);
                    stage_store(&tm_yh, c0, reinterpret_cast<const uint4 *>(h), b * p.H + y, xs);
                    stage_store(&tm_yl, c0, reinterpret_cast<const uint4 *>(l), b * p.H + y, xs);
                }
                if (p.has_f32) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] *= 1.f / kScale;
                    stage_store(&tm_y32, c0, reinterpret_cast<const uint4 *>(v), b * p.H + y, xs);
                    if (c0 + 16 < p.y32_stride) stage_store(&tm_y32, c0 + 16, reinterpret_cast<const uint4 *>(v + 16), b * p.H + y, xs);
                }
            }
        }
        if (lane == 0) tc::tma_store_wait_all();
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    if (warp == 9) tc::tmem_dealloc<Cfg<BN>::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------------------------- 3x3 with halo re-use
// The same GEMM with a third of the activation traffic: the tile is 8 pixels wide and 16 rows tall (GEMM row m = 8 * row + column), so
// one image row of the tile is exactly one 1024-byte swizzle atom and a shift by one image row is a shift of the operand's start
// address by one atom.  Per (kx, 64-channel block) ONE halo box {64 ch, 8 px, 18 rows} is loaded and feeds the three ky taps as
// descriptors at +0, +1 KB, +2 KB.  Activations and weights travel in separate rings (a halo box lives three times as long as a
// weight tap).  With POOL the epilogue takes the 2x2 maximum across lanes (m ^ 1: neighbour in x, m ^ 8: neighbour in y) and
// stores the pooled tile -- MaxPool2d(2,2) never touches memory.
template <int BN>
struct HaloCfg {
    static constexpr int A_BUF = 2 * 18 * 1024;                       // hi + lo halo boxes
    static constexpr int W_BUF = 2 * BN * BK * 2;                     // hi + lo of one tap
    static constexpr int NA = BN == 64 ? 3 : 2;
    static constexpr int NW = BN == 64 ? 5 : 3;
    static constexpr int SMEM_BYTES = NA * A_BUF + NW * W_BUF + STG_BYTES + BIAS_FLOATS * 4 + 256;
    // BN = 64: an M128 x N64 x K16 MMA occupies the tensor pipe as long as an N128 one (measured ~65-90 clk: the 4 KB A operand is the
    // cost), so Ah*Wh and Ah*Wl are issued as ONE N = 128 MMA against the adjacent [Wh; Wl] rows of the weight buffer: two MMAs per
    // k-step instead of three, accumulator = 128 columns (hi*hi + lo*hi | hi*lo), added in the epilogue.
    static constexpr bool WIDE = BN == 64;
    static constexpr int ACC_COLS = WIDE ? 2 * BN : BN;
    static constexpr int TMEM_COLS = 2 * ACC_COLS;
};

struct HaloBarriers {
    uint64_t a_full[3], a_empty[3], w_full[5], w_empty[5], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

template <int BN, int POOL>
__global__ void __launch_bounds__(THREADS, 1)
k_conv_halo(const __grid_constant__ CUtensorMap tm_xh, const __grid_constant__ CUtensorMap tm_xl, const __grid_constant__ CUtensorMap tm_wh,
            const __grid_constant__ CUtensorMap tm_wl, const __grid_constant__ CUtensorMap tm_yh, const __grid_constant__ CUtensorMap tm_yl,
            const Params p) {
    using C = HaloCfg<BN>;
    constexpr int NA = C::NA, NW = C::NW, A_BUF = C::A_BUF, W_BUF = C::W_BUF, A_HALF = A_BUF / 2, W_HALF = W_BUF / 2;
    extern __shared__ __align__(1024) uint8_t smem[];
    if (tc::smem_u32(smem) & 1023u) __trap();
    uint8_t *s_a = smem, *s_w = smem + NA * A_BUF, *s_stg = s_w + NW * W_BUF;
    float *s_bias = reinterpret_cast<float *>(s_stg + STG_BYTES);
    HaloBarriers *bar = reinterpret_cast<HaloBarriers *>(s_stg + STG_BYTES + BIAS_FLOATS * 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkb = p.Cin / BK;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NA; ++i) { tc::mbar_init(&bar->a_full[i], 1); tc::mbar_init(&bar->a_empty[i], 1); }
        for (int i = 0; i < NW; ++i) { tc::mbar_init(&bar->w_full[i], 1); tc::mbar_init(&bar->w_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&bar->acc_full[i], 1); tc::mbar_init(&bar->acc_empty[i], EPI_WARPS); }
        tc::fence_barrier_init();
    }
    for (int i = threadIdx.x; i < BIAS_FLOATS; i += THREADS) s_bias[i] = (p.bias && i < p.Cout) ? kScale * p.bias[i] : 0.f;
    if (warp == 9) tc::tmem_alloc<C::TMEM_COLS>(&bar->tmem_base);
    if (warp == 8 && lane == 0) {
        tc::tma_prefetch_desc(&tm_xh); tc::tma_prefetch_desc(&tm_xl); tc::tma_prefetch_desc(&tm_wh); tc::tma_prefetch_desc(&tm_wl);
        tc::tma_prefetch_desc(&tm_yh); tc::tma_prefetch_desc(&tm_yl);
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = bar->tmem_base;

    auto decode = [&](int t, int &b, int &y0, int &x0, int &ct) {
        ct = t % p.col_tiles; int r = t / p.col_tiles;
        x0 = (r % p.tiles_x) * 8; r /= p.tiles_x;
        y0 = (r % p.tiles_y) * 16; b = r / p.tiles_y;
    };

    if (warp == 8) {
        // ================= TMA producer: A(kx, kb) then its three weight taps, in the order the MMA warp consumes them =================
        if (lane == 0 && !(p.dbg & 1)) {
            uint32_t ia = 0, iw = 0;
            for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
                int b, y0, x0, ct; decode(t, b, y0, x0, ct);
                for (int kb = 0; kb < nkb; ++kb)
                    for (int kx = 0; kx < 3; ++kx) {
                        {
                            const uint32_t s = ia % NA, ph = (ia / NA) & 1;
                            uint8_t *st = s_a + s * A_BUF;
                            tc::mbar_wait(&bar->a_empty[s], ph ^ 1);
                            tc::mbar_arrive_expect_tx(&bar->a_full[s], A_BUF);
                            tc::tma_load_4d(st, &tm_xh, &bar->a_full[s], kb * BK, x0 + kx - 1, y0 - 1, b);
                            tc::tma_load_4d(st + A_HALF, &tm_xl, &bar->a_full[s], kb * BK, x0 + kx - 1, y0 - 1, b);
                            ++ia;
                        }
                        for (int ky = 0; ky < 3; ++ky, ++iw) {
                            const uint32_t s = iw % NW, ph = (iw / NW) & 1;
                            uint8_t *st = s_w + s * W_BUF;
                            tc::mbar_wait(&bar->w_empty[s], ph ^ 1);
                            tc::mbar_arrive_expect_tx(&bar->w_full[s], W_BUF);
                            tc::tma_load_3d(st, &tm_wh, &bar->w_full[s], kb * BK, ct * BN, ky * 3 + kx);
                            tc::tma_load_3d(st + W_HALF, &tm_wl, &bar->w_full[s], kb * BK, ct * BN, ky * 3 + kx);
                        }
                    }
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = tc::umma_idesc_f16(BM, BN), idesc_wide = tc::umma_idesc_f16(BM, 2 * BN);
            const uint32_t a_base = tc::smem_u32(s_a), w_base = tc::smem_u32(s_w);
            uint32_t ia = 0, iw = 0;
            int j = 0;
            for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++j) {
                const uint32_t as = j & 1, aph = (j >> 1) & 1;
                tc::mbar_wait(&bar->acc_empty[as], aph ^ 1);
                tc::tcgen05_fence_after();
                const uint32_t d = tmem_base + as * C::ACC_COLS;
                uint32_t first = 0;                                      // 0 until the tile's first MMA has been issued
                for (int kk = 0; kk < 3 * nkb; ++kk, ++ia) {
                    const uint32_t sa = ia % NA, pha = (ia / NA) & 1;
                    if (!(p.dbg & 1)) tc::mbar_wait(&bar->a_full[sa], pha);
                    tc::tcgen05_fence_after();
                    for (int ky = 0; ky < 3; ++ky, ++iw) {
                        const uint32_t sw = iw % NW, phw = (iw / NW) & 1;
                        const uint32_t ab = a_base + sa * A_BUF + ky * 1024, wb = w_base + sw * W_BUF;
                        const uint64_t ah = tc::umma_desc_k128(ab), al = tc::umma_desc_k128(ab + A_HALF);
                        const uint64_t wh = tc::umma_desc_k128(wb), wl = tc::umma_desc_k128(wb + W_HALF);
                        if (!(p.dbg & 1)) tc::mbar_wait(&bar->w_full[sw], phw);
                        tc::tcgen05_fence_after();
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            if (p.dbg & 16) continue;
                            if (C::WIDE) {
                                tc::umma_f16(d, ah + 2 * k, wh + 2 * k, idesc_wide, first);     // Ah * [Wh; Wl] -> columns 0..63 | 64..127
                                first = 1;
                                if (p.dbg & 2) continue;
                                tc::umma_f16(d, al + 2 * k, wh + 2 * k, idesc, 1);              // Al * Wh -> columns 0..63
                                continue;
                            }
                            tc::umma_f16(d, ah + 2 * k, wh + 2 * k, idesc, first);
                            first = 1;
                            if (p.dbg & 2) continue;
                            tc::umma_f16(d, al + 2 * k, wh + 2 * k, idesc, 1);
                            tc::umma_f16(d, ah + 2 * k, wl + 2 * k, idesc, 1);
                        }
                        tc::umma_commit(&bar->w_empty[sw]);
                    }
                    tc::umma_commit(&bar->a_empty[sa]);
                }
                tc::umma_commit(&bar->acc_full[as]);
            }
        }
        __syncwarp();
    } else {
        // ================= epilogue: warp w owns GEMM rows 32 * (w % 4) .., i.e. image rows 4 * (w % 4) .. + 3 of the tile, 8 pixels each =================
        const int quad = warp & 3, half = warp >> 2;
        const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
        uint32_t nst = 0;
        // staging row r holds 64 bytes of one pixel; the TMA box is {32 channels, 8 px, 4 rows} (or {32, 4, 2} pooled)
        auto stage_store = [&](const CUtensorMap *tm, int col, const uint4 *data, bool writer, int r, int x, int y, int b) {
            if (p.dbg & 4) return;
            uint8_t *buf = s_stg + (warp * 2 + (nst & 1)) * STG_BUF;
            if (lane == 0) tc::tma_store_wait_read1();
            __syncwarp();
            if (writer) {
#pragma unroll
                for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4 *>(buf + r * 64 + ((q ^ ((r >> 1) & 3)) << 4)) = data[q];
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) { tc::tma_store_4d(tm, buf, col, x, y, b); tc::tma_store_commit(); }
            ++nst;
        };
        constexpr int CHUNKS = BN / 64;
        int j = 0;
        for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++j) {
            int b, y0, x0, ct; decode(t, b, y0, x0, ct);
            const uint32_t as = j & 1, aph = (j >> 1) & 1;
            tc::mbar_wait(&bar->acc_full[as], aph);
            tc::tcgen05_fence_after();
#pragma unroll 1
            for (int ch = 0; ch < CHUNKS; ++ch) {
                const int ccol = half * (BN / 2) + ch * 32;
                const int c0 = ct * BN + ccol;
                float v[32];
                if (!(p.dbg & 32)) {
                    tc::tmem_ld_32x32(lane_base + as * C::ACC_COLS + ccol, v);
                    if (C::WIDE) {
                        float u[32];
                        tc::tmem_ld_32x32(lane_base + as * C::ACC_COLS + BN + ccol, u);
                        tc::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] += u[i];
                    } else {
                        tc::tmem_ld_wait();
                    }
                }
                if (ch == CHUNKS - 1) {
                    tc::tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(&bar->acc_empty[as]);
                }
                if (c0 >= p.Cout || y0 + 4 * quad >= p.H || (p.dbg & 8)) continue;            // padding channels / rows below the image (warp-uniform)
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float4 bq = *reinterpret_cast<const float4 *>(s_bias + c0 + 4 * q);
                    v[4 * q] = fmaf(v[4 * q], 1.f / kScale, bq.x); v[4 * q + 1] = fmaf(v[4 * q + 1], 1.f / kScale, bq.y);
                    v[4 * q + 2] = fmaf(v[4 * q + 2], 1.f / kScale, bq.z); v[4 * q + 3] = fmaf(v[4 * q + 3], 1.f / kScale, bq.w);
                }
                if (p.relu) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
                }
                uint32_t h[16], l[16];
                if (POOL) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        v[i] = fmaxf(v[i], __shfl_xor_sync(0xffffffffu, v[i], 1));
                        v[i] = fmaxf(v[i], __shfl_xor_sync(0xffffffffu, v[i], 8));
                    }
                    split_planes32(v, h, l);
                    const bool writer = (lane & 9) == 0;                          // even column, even row of the 8 x 4 patch
                    const int r = (lane >> 4) * 4 + ((lane & 7) >> 1);            // pooled pixel: row lane / 16, column (lane % 8) / 2
                    stage_store(&tm_yh, c0, reinterpret_cast<const uint4 *>(h), writer, r, x0 >> 1, (y0 >> 1) + 2 * quad, b);
                    stage_store(&tm_yl, c0, reinterpret_cast<const uint4 *>(l), writer, r, x0 >> 1, (y0 >> 1) + 2 * quad, b);
                } else {
                    split_planes32(v, h, l);
                    stage_store(&tm_yh, c0, reinterpret_cast<const uint4 *>(h), true, lane, x0, y0 + 4 * quad, b);
                    stage_store(&tm_yl, c0, reinterpret_cast<const uint4 *>(l), true, lane, x0, y0 + 4 * quad, b);
                }
            }
        }
        if (lane == 0) tc::tma_store_wait_all();
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    if (warp == 9) tc::tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

template <int BN, int POOL>
static int launch_halo(const CUtensorMap *maps, const Params &p, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        OPB_CHECK_CUDA(cudaFuncSetAttribute(k_conv_halo<BN, POOL>, cudaFuncAttributeMaxDynamicSharedMemorySize, HaloCfg<BN>::SMEM_BYTES));
        configured = true;
    }
    int dev = 0, sms = 0;
    OPB_CHECK_CUDA(cudaGetDevice(&dev));
    OPB_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = p.total_tiles < sms ? p.total_tiles : sms;
    k_conv_halo<BN, POOL><<<grid, THREADS, HaloCfg<BN>::SMEM_BYTES, st>>>(maps[0], maps[1], maps[2], maps[3], maps[4], maps[5], p);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

// ---------------------------------------------------------------------------------------------- first layer: 1 -> 64 channels, 3x3
// Block = 32 x 8 pixels.  Thread = (pixel slot tid / 8, channel group tid % 8): its 8 x 9 weights live in registers for the whole
// block, the 8 lanes of a pixel write 128 contiguous bytes per plane and the 4 pixels of a warp are neighbours in x, so every
// warp store is one contiguous 512-byte run of the NHWC planes.
__global__ void __launch_bounds__(256)
k_conv1a_planes(const float *__restrict__ img, const float *__restrict__ w, const float *__restrict__ bias, int H, int W,
                __half *__restrict__ yh, __half *__restrict__ yl) {
    __shared__ float s_in[10][34];
    const int b = blockIdx.z, x0 = blockIdx.x * 32, y0 = blockIdx.y * 8, tid = threadIdx.x;
    const float *im = img + (long long)b * H * W;
    for (int i = tid; i < 340; i += 256) {
        const int ty = i / 34, tx = i % 34, y = y0 + ty - 1, x = x0 + tx - 1;
        s_in[ty][tx] = (y >= 0 && y < H && x >= 0 && x < W) ? im[(long long)y * W + x] : 0.f;
    }
    const int g = tid & 7, slot = tid >> 3;                    // channels 8g .. 8g+7; pixels slot, slot + 32, ...
    float wr[8][9], br[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        br[c] = bias[g * 8 + c];
#pragma unroll
        for (int k = 0; k < 9; ++k) wr[c][k] = w[(g * 8 + c) * 9 + k];
    }
    __syncthreads();
#pragma unroll 1
    for (int pi = slot; pi < 256; pi += 32) {
        const int tx = pi & 31, ty = pi >> 5, x = x0 + tx, y = y0 + ty;
        if (x >= W || y >= H) continue;
        float in[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) in[k] = s_in[ty + k / 3][tx + k % 3];
        float v[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < 9; ++k) acc = fmaf(wr[c][k], in[k], acc);
            v[c] = fmaxf(acc + br[c], 0.f) * kScale;
        }
        uint32_t h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float a0 = fminf(v[2 * i], 65000.f), a1 = fminf(v[2 * i + 1], 65000.f);
            const __half2 hi = __floats2half2_rn(a0, a1);
            const float2 hf = __half22float2(hi);
            const __half2 lo = __floats2half2_rn(a0 - hf.x, a1 - hf.y);
            h[i] = *reinterpret_cast<const uint32_t *>(&hi); l[i] = *reinterpret_cast<const uint32_t *>(&lo);
        }
        const long long o = (((long long)b * H + y) * W + x) * 64 + g * 8;
        *reinterpret_cast<uint4 *>(yh + o) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4 *>(yl + o) = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

// ---------------------------------------------------------------------------------------------- 2x2 max-pool on planes
// thread = (output pixel, group of 8 channels)
__global__ void __launch_bounds__(256)
k_pool2_planes(const __half *__restrict__ xh, const __half *__restrict__ xl, int B, int H, int W, int C, __half *__restrict__ yh,
               __half *__restrict__ yl) {
    const int Ho = H / 2, Wo = W / 2, G = C / 8;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * Ho * Wo * G) return;
    const int g = (int)(i % G);
    long long r = i / G;
    const int xo = (int)(r % Wo); r /= Wo;
    const int yo = (int)(r % Ho);
    const int b = (int)(r / Ho);
    float best[8];
    uint4 bh = make_uint4(0, 0, 0, 0), bl = bh;
    __half *ph = reinterpret_cast<__half *>(&bh), *pl = reinterpret_cast<__half *>(&bl);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const long long src = ((((long long)b * H + 2 * yo + (k >> 1)) * W + 2 * xo + (k & 1)) * C) + g * 8;
        const uint4 h4 = *reinterpret_cast<const uint4 *>(xh + src), l4 = *reinterpret_cast<const uint4 *>(xl + src);
        const __half *hh = reinterpret_cast<const __half *>(&h4), *ll = reinterpret_cast<const __half *>(&l4);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const float v = __half2float(hh[c]) + __half2float(ll[c]);
            if (k == 0 || v > best[c]) { best[c] = v; ph[c] = hh[c]; pl[c] = ll[c]; }
        }
    }
    const long long dst = ((((long long)b * Ho + yo) * Wo + xo) * C) + g * 8;
    *reinterpret_cast<uint4 *>(yh + dst) = bh;
    *reinterpret_cast<uint4 *>(yl + dst) = bl;
}

template <int BN>
static int launch_conv(const CUtensorMap *maps, const Params &p, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        OPB_CHECK_CUDA(cudaFuncSetAttribute(k_conv_tc<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
        configured = true;
    }
    int dev = 0, sms = 0;
    OPB_CHECK_CUDA(cudaGetDevice(&dev));
    OPB_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = p.total_tiles < sms ? p.total_tiles : sms;
    k_conv_tc<BN><<<grid, THREADS, Cfg<BN>::SMEM_BYTES, st>>>(maps[0], maps[1], maps[2], maps[3], maps[4], maps[5], maps[6], p);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

}  // namespace convtc
}  // namespace opb

extern "C" int op_conv_tc(const void *xh, const void *xl, int B, int H, int W, int Cin, const void *wh, const void *wl, int Cout_pad,
                          const float *bias, int Cout, int ntaps, int relu, int pool, void *yh, void *yl, float *y32, int y32_stride,
                          void *stream) {
    using namespace opb;
    using namespace opb::convtc;
    OPB_CHECK_ARG(xh && xl && wh && wl && B > 0 && H > 0 && W > 0, "op_conv_tc: bad argument");
    OPB_CHECK_ARG(Cin >= 64 && Cin % 64 == 0, "op_conv_tc: Cin=%d must be a multiple of 64", Cin);
    OPB_CHECK_ARG(ntaps == 9 || ntaps == 1, "op_conv_tc: ntaps must be 9 (3x3, pad 1) or 1 (1x1)");
    OPB_CHECK_ARG(Cout > 0 && Cout <= Cout_pad && Cout_pad <= BIAS_FLOATS && (Cout_pad == 64 || Cout_pad % 128 == 0),
                  "op_conv_tc: Cout=%d / padded %d (64 or a multiple of 128, at most 256)", Cout, Cout_pad);
    OPB_CHECK_ARG((yh && yl) || y32, "op_conv_tc: no output");
    OPB_CHECK_ARG(!(yh && yl) || Cout % 32 == 0, "op_conv_tc: plane output needs Cout %% 32 == 0");
    OPB_CHECK_ARG(!y32 || (y32_stride >= Cout && y32_stride % 4 == 0), "op_conv_tc: y32_stride=%d", y32_stride);
    Params p{};
    p.B = B; p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout; p.ntaps = ntaps; p.relu = relu;
    p.bias = bias;
    static const int use_halo = [] { const char *e = getenv("OPB_CONV_HALO"); return e ? atoi(e) : 1; }();
    static const int dbg = [] { const char *e = getenv("OPB_CONV_DBG"); return e ? atoi(e) : 0; }();
    p.dbg = dbg;
    const bool halo = use_halo && ntaps == 9 && yh && yl && !y32 && W % 8 == 0;
    OPB_CHECK_ARG(!pool || (halo && H % 2 == 0), "op_conv_tc: the fused 2x2 max-pool needs a 3x3 convolution with plane output, W %% 8 == 0 and even H");
    if (halo) {
        const int BN = Cout_pad == 64 ? 64 : 128;
        p.TW = 8; p.TH = 16; p.tiles_x = W / 8; p.tiles_y = ceil_div(H, 16); p.col_tiles = Cout_pad / BN;
        const long long total = (long long)B * p.tiles_x * p.tiles_y * p.col_tiles;
        OPB_CHECK_ARG(total < (1ll << 31), "op_conv_tc: too many tiles");
        p.total_tiles = (int)total; p.has_planes = 1;
        CUtensorMap maps[6];
        int rc;
        if ((rc = make_tmap_f16_nhwc(&maps[0], xh, B, H, W, Cin, 8, 18)) != OPB_OK) return rc;
        if ((rc = make_tmap_f16_nhwc(&maps[1], xl, B, H, W, Cin, 8, 18)) != OPB_OK) return rc;
        if ((rc = make_tmap_f16_3d(&maps[2], wh, 9, Cout_pad, Cin, BN)) != OPB_OK) return rc;
        if ((rc = make_tmap_f16_3d(&maps[3], wl, 9, Cout_pad, Cin, BN)) != OPB_OK) return rc;
        const int Ho = pool ? H / 2 : H, Wo = pool ? W / 2 : W;
        if ((rc = make_tmap_f16_nhwc_store(&maps[4], yh, B, Ho, Wo, Cout, pool ? 4 : 8, pool ? 2 : 4)) != OPB_OK) return rc;
        if ((rc = make_tmap_f16_nhwc_store(&maps[5], yl, B, Ho, Wo, Cout, pool ? 4 : 8, pool ? 2 : 4)) != OPB_OK) return rc;
        cudaStream_t st = (cudaStream_t)stream;
        if (BN == 64) return pool ? launch_halo<64, 1>(maps, p, st) : launch_halo<64, 0>(maps, p, st);
        return pool ? launch_halo<128, 1>(maps, p, st) : launch_halo<128, 0>(maps, p, st);
    }
    p.TW = W >= 128 ? 128 : (W >= 64 ? 64 : 32); p.TH = BM / p.TW;
    p.tiles_x = ceil_div(W, p.TW); p.tiles_y = ceil_div(H, p.TH);
    const int BN = Cout_pad == 64 ? 64 : 128;
    p.col_tiles = Cout_pad / BN;
    const long long total = (long long)B * p.tiles_x * p.tiles_y * p.col_tiles;
    OPB_CHECK_ARG(total < (1ll << 31), "op_conv_tc: too many tiles");
    p.total_tiles = (int)total;
    p.y32_stride = y32_stride; p.has_planes = (yh && yl) ? 1 : 0; p.has_f32 = y32 ? 1 : 0;
    CUtensorMap maps[7];
    int rc;
    if ((rc = make_tmap_f16_nhwc(&maps[0], xh, B, H, W, Cin, p.TW, p.TH)) != OPB_OK) return rc;
    if ((rc = make_tmap_f16_nhwc(&maps[1], xl, B, H, W, Cin, p.TW, p.TH)) != OPB_OK) return rc;
    if ((rc = make_tmap_f16_3d(&maps[2], wh, ntaps, Cout_pad, Cin, BN)) != OPB_OK) return rc;
    if ((rc = make_tmap_f16_3d(&maps[3], wl, ntaps, Cout_pad, Cin, BN)) != OPB_OK) return rc;
    maps[4] = maps[0]; maps[5] = maps[0]; maps[6] = maps[0];
    if (p.has_planes) {
        if ((rc = make_tmap_f16_3d_store(&maps[4], yh, B * H, W, Cout, 32)) != OPB_OK) return rc;
        if ((rc = make_tmap_f16_3d_store(&maps[5], yl, B * H, W, Cout, 32)) != OPB_OK) return rc;
    }
    if (p.has_f32 && (rc = make_tmap_f32_3d_store(&maps[6], y32, B * H, W, y32_stride, 32)) != OPB_OK) return rc;
    return BN == 64 ? launch_conv<64>(maps, p, (cudaStream_t)stream) : launch_conv<128>(maps, p, (cudaStream_t)stream);
}

extern "C" int op_conv1a_planes(const float *img, const float *w, const float *bias, int B, int H, int W, void *yh, void *yl, void *stream) {
    OPB_CHECK_ARG(img && w && bias && yh && yl && B > 0 && H > 0 && W > 0 && B <= 65535, "op_conv1a_planes: bad argument");
    opb::convtc::k_conv1a_planes<<<dim3(opb::ceil_div(W, 32), opb::ceil_div(H, 8), B), 256, 0, (cudaStream_t)stream>>>(
        img, w, bias, H, W, (__half *)yh, (__half *)yl);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_pool2_planes(const void *xh, const void *xl, int B, int H, int W, int C, void *yh, void *yl, void *stream) {
    OPB_CHECK_ARG(xh && xl && yh && yl && B > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0 && C % 8 == 0, "op_pool2_planes: bad argument");
    const long long n = (long long)B * (H / 2) * (W / 2) * (C / 8);
    opb::convtc::k_pool2_planes<<<(unsigned)opb::ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>((const __half *)xh, (const __half *)xl, B, H, W, C,
                                                                                                   (__half *)yh, (__half *)yl);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
