// tcgen05 / TMA version of the similarity + dual-softmax passes (included by sim_dual_softmax.cu).
//
// A CTA owns 128 query rows of one frame; by default two CTAs form a pair (cta_group::2) that walks 256-column tiles of the 3D
// points with M256 x N256 MMAs (CG = 2); CG = 1 is the one-CTA form on 128-column tiles.
//   warp 8   : TMA producer.  The CTA's A operand (128 rows x K=256, 128 KB of planes) is loaded once and stays resident; B (3D
//              descriptors) streams through a ring of four 16 KB units (a unit = 128 columns x 128 bytes of one plane; in a pair
//              each CTA stages its half of the tile's columns and both CTAs' bytes are credited to the leader's barrier).
//   warp 9   : TMEM allocation + single-thread tcgen05.mma issue (the pair's leader only).  fp32 accumulators are double-buffered in
//              TMEM (2 x 128 / 2 x 256 columns) so the epilogue of tile j overlaps the MMAs of tile j+1.
//   warps 0-7: epilogue.  Thread = one query row (TMEM lane), warp = (lane quadrant, column half).  e = exp2(acc*c1 - c0) with the
//              fixed shift 1/scale (|cos| <= 1).  PASS 1 keeps the row sum in a register and reduces column sums across the 128
//              rows with a shuffle transpose-reduce; PASS 2 writes conf = (e/R)(e/C) once through swizzled staging + TMA stores and
//              tracks row maxima (16-column blocks) and column maxima (transposed read of the staged block).
#pragma once
#include <cuda_fp8.h>
#include "tc_common.cuh"

namespace opb {
namespace simtc {

constexpr int BM = 128, NKB = 4;
constexpr int TILE_BYTES = 128 * 128;                 // 128 rows x 64 fp16
constexpr int A_BYTES = 2 * NKB * TILE_BYTES;         // 131072
constexpr int RING_BYTES = 4 * TILE_BYTES;            // 65536
constexpr int MAX_UNITS = 4;
// CG = 1: one CTA per 128 query rows, M128 x N128 MMAs on 128-column tiles (OPB_SIM_CG=1, A/B runs only).
// CG = 2 (default since r01zb): a CTA pair (cta_group::2) per 256 query rows on 256-column tiles.  Each CTA keeps its own 128 rows
//         of A resident and stages HALF of every B k-block in 16 KB ring units (128 columns x 128 bytes of one plane); the leader
//         issues M256 x N256 MMAs, which halves the L2 -> SM operand stream per flop that bounds the one-CTA form.
// The production kernels are the sixteen-epilogue-warp pair kernels of sim_tc16.cuh (k_sim16); the eight-warp kernels in this file
// remain as the A/B forms (OPB_SIM_EW=8, OPB_SIM_CG=1, OPB_SIM_MMA=f16) and as the source of the shared constants / helpers.
template <int CG>
struct SimCfg {
    static constexpr int BN = 128 * CG;               // columns of a tile: 128 (one CTA) / 256 (pair, M256 x N256 like the projections)
    static constexpr int UNITS = 4;
    static constexpr int B_ROWS = 128;                // columns of a B k-block held by one CTA
    static constexpr int UNIT_BYTES = B_ROWS * 128;   // 16384
    static constexpr int HALF = BN / 2;               // columns of a tile one epilogue warp owns: 64 / 128
    static constexpr int GROUPS = HALF / 64;          // pairs of 32-column chunks per warp and tile
    static constexpr int TMEM_COLS = 2 * BN;          // double-buffered accumulator: 256 / 512 columns
};
constexpr int STG_WARP_BYTES = 32 * 16 * 4;            // per epilogue warp: 32 rows x 16 fp32 columns, TMA-store staging
constexpr int STG_BYTES = 8 * STG_WARP_BYTES;         // 16384
constexpr int SCRATCH_BYTES = 2 * 4 * 256 * 4;        // double-buffered [4 quadrants][<= 256 columns] of 32-bit partials
constexpr int LGC_BYTES = 2 * 256 * 4;                // double-buffered column terms of the current tile
constexpr int SMEM_BYTES = A_BYTES + RING_BYTES + STG_BYTES + SCRATCH_BYTES + LGC_BYTES + 256 + 1024;   // + barriers + alignment slack
constexpr float kOperandScale = 64.f;                 // unit-vector components are scaled into fp16's normal range

struct Params {
    int N2, N3, tiles_n, tiles_per_split;
    float c1, c0;                                     // e = exp2(acc * c1 - c0)
    float *rowsum_part, *colsum_part;                 // [B][tiles_n][N2] (per column tile), [B][tiles_m][N3]
    float c2;                                         // 2*c1
    const float *lgR, *lgC;                           // log2(1/R_n), log2(1/C_m)
    unsigned long long *rowmax_part, *colmax_part;
    float *conf;
    int tiles_m;                                      // row tiles of a frame (the grid may carry one ghost row tile to fill a pair)
    int tma_store;                                    // conf rows are 16-byte aligned: write conf through TMA stores
    int dbg;                                          // experiment switches (OPB_SIM_DBG; timing only): 1 = no conf stores, 8 = no MMAs; 0 in production
};

struct Barriers {
    uint64_t a_full, b_full[MAX_UNITS], b_empty[MAX_UNITS], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// sum over the 32 lanes of v[i]; lane L ends up with the total of element L in v[0]
__device__ __forceinline__ float transpose_reduce_sum(float *v, int lane) {
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        const bool upper = (lane & step) != 0;
#pragma unroll
        for (int i = 0; i < step; ++i) {
            const float send = upper ? v[i] : v[i + step];
            const float keep = upper ? v[i + step] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, step);
        }
    }
    return v[0];
}
__device__ __forceinline__ unsigned long long transpose_reduce_max(unsigned long long *v, int lane) {
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        const bool upper = (lane & step) != 0;
#pragma unroll
        for (int i = 0; i < step; ++i) {
            const unsigned long long send = upper ? v[i] : v[i + step];
            const unsigned long long keep = upper ? v[i + step] : v[i];
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, send, step);
            v[i] = keep > other ? keep : other;
        }
    }
    return v[0];
}

// F8 = 0: a product is three kind::f16 MMAs, hi*hi + lo*hi + hi*lo on fp16 planes (tm_*l = the lo planes, tm_*x unused).
// F8 = 1 (pair form only): hi*hi on the fp16 hi planes plus the two cross terms as kind::f8f6f4 MMAs at twice the rate,
//         lo(e5m2) * hi(e4m3) + hi(e4m3) * lo(e5m2) on byte planes (tm_*l = lo as e5m2, tm_*x = hi as e4m3, [B, N, 256] bytes).  The cross
//         terms are 2^-11 of the product, so 3 mantissa bits keep the cosine to 5.4e-6 (1.6e-4 on conf; CPU emulation and
//         tests/test_gpu_matcher.py) at two thirds of the tensor work and the same operand bytes.
template <int PASS, int CG, int F8>
__global__ void __launch_bounds__(320, 1)
k_sim_tc(const __grid_constant__ CUtensorMap tm_m2h, const __grid_constant__ CUtensorMap tm_m2l, const __grid_constant__ CUtensorMap tm_m2x,
         const __grid_constant__ CUtensorMap tm_m3h, const __grid_constant__ CUtensorMap tm_m3l, const __grid_constant__ CUtensorMap tm_m3x,
         const __grid_constant__ CUtensorMap tm_conf, const Params p) {
    static_assert(!F8 || CG == 2, "the fp8 cross terms are implemented for the pair form");
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment (SWIZZLE_128B) by an offset on the shared array, so accesses stay in the shared address space (LDS/STS)
    uint8_t *smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *sA = smem, *sB = smem + A_BYTES;
    uint8_t *s_stg = smem + A_BYTES + RING_BYTES;
    uint32_t *scratch = reinterpret_cast<uint32_t *>(smem + A_BYTES + RING_BYTES + STG_BYTES);
    float *s_lgc = reinterpret_cast<float *>(smem + A_BYTES + RING_BYTES + STG_BYTES + SCRATCH_BYTES);
    Barriers *bar = reinterpret_cast<Barriers *>(smem + A_BYTES + RING_BYTES + STG_BYTES + SCRATCH_BYTES + LGC_BYTES);

    using Cfg = SimCfg<CG>;
    constexpr int BN = Cfg::BN, UNITS = Cfg::UNITS, UNIT_BYTES = Cfg::UNIT_BYTES, HALF = Cfg::HALF, GROUPS = Cfg::GROUPS;
    const uint32_t rank = CG == 2 ? tc::cluster_ctarank() : 0;            // 0 = leader of the pair (cluster dims (2, 1, 1): row tiles 2i, 2i+1)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int split = blockIdx.y, row_tile = blockIdx.x, b = blockIdx.z;
    const bool tile_ok = row_tile < p.tiles_m;                            // false: ghost row tile completing a pair
    const int m0 = row_tile * BM;
    const int t0 = split * p.tiles_per_split;
    const int t1 = min(p.tiles_n, t0 + p.tiles_per_split);
    if (t0 >= t1) return;                                   // uniform for the whole CTA (and its pair)

    if (threadIdx.x == 0) {
        tc::mbar_init(&bar->a_full, 1);
        for (int i = 0; i < UNITS; ++i) { tc::mbar_init(&bar->b_full[i], 1); tc::mbar_init(&bar->b_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&bar->acc_full[i], 1); tc::mbar_init(&bar->acc_empty[i], 8 * CG); }
        tc::fence_barrier_init();
    }
    if (warp == 9) { if (CG == 2) tc::tmem_alloc_cg2<Cfg::TMEM_COLS>(&bar->tmem_base); else tc::tmem_alloc<Cfg::TMEM_COLS>(&bar->tmem_base); }
    if (warp == 8 && lane == 0) {
        tc::tma_prefetch_desc(&tm_m2h); tc::tma_prefetch_desc(&tm_m2l); tc::tma_prefetch_desc(&tm_m3h); tc::tma_prefetch_desc(&tm_m3l);
        if (F8) { tc::tma_prefetch_desc(&tm_m2x); tc::tma_prefetch_desc(&tm_m3x); }
        if (PASS == 2 && p.tma_store) tc::tma_prefetch_desc(&tm_conf);
    }
    tc::tcgen05_fence_before();
    if (CG == 2) tc::cluster_sync(); else __syncthreads();                // the peer's barriers exist before anything is sent to them
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = bar->tmem_base;

    if (warp == 8) {
        // ================= TMA producer =================
        if (lane == 0 && F8) {
            // A resident: hi fp16 (4 k-blocks of 64, 64 KB), then lo e5m2 and hi e4m3 (2 k-blocks of 128 bytes each, 32 + 32 KB).
            // B ring, 8 units of 16 KB per tile: 4 x hi fp16, 2 x hi e4m3, 2 x lo e5m2
            const uint32_t af = tc::mapa(tc::smem_u32(&bar->a_full), 0);
            if (rank == 0) tc::mbar_arrive_expect_tx(&bar->a_full, 2 * A_BYTES);
            for (int kb = 0; kb < NKB; ++kb) tc::tma_load_3d_cg2(sA + kb * TILE_BYTES, &tm_m2h, af, kb * 64, m0, b);
            for (int j = 0; j < 2; ++j) tc::tma_load_3d_cg2(sA + (4 + j) * TILE_BYTES, &tm_m2l, af, j * 128, m0, b);
            for (int j = 0; j < 2; ++j) tc::tma_load_3d_cg2(sA + (6 + j) * TILE_BYTES, &tm_m2x, af, j * 128, m0, b);
            uint32_t u = 0;
            for (int t = t0; t < t1; ++t)
                for (int q = 0; q < 8; ++q, ++u) {
                    const uint32_t stage = u % UNITS, ph = (u / UNITS) & 1;
                    tc::mbar_wait(&bar->b_empty[stage], ph ^ 1);
                    if (rank == 0) tc::mbar_arrive_expect_tx(&bar->b_full[stage], 2 * UNIT_BYTES);
                    const CUtensorMap *tm = q < 4 ? &tm_m3h : (q < 6 ? &tm_m3x : &tm_m3l);
                    const int c0 = q < 4 ? q * 64 : (q & 1) * 128;
                    tc::tma_load_3d_cg2(sB + stage * UNIT_BYTES, tm, tc::mapa(tc::smem_u32(&bar->b_full[stage]), 0), c0,
                                        t * BN + (int)rank * Cfg::B_ROWS, b);
                }
        }
        if (lane == 0 && !F8) {
            if (CG == 2) {
                // both CTAs load into their own shared memory; the bytes of both are credited to the leader's barriers
                const uint32_t af = tc::mapa(tc::smem_u32(&bar->a_full), 0);
                if (rank == 0) tc::mbar_arrive_expect_tx(&bar->a_full, 2 * A_BYTES);
                for (int plane = 0; plane < 2; ++plane)
                    for (int kb = 0; kb < NKB; ++kb)
                        tc::tma_load_3d_cg2(sA + (plane * NKB + kb) * TILE_BYTES, plane ? &tm_m2l : &tm_m2h, af, kb * 64, m0, b);
            } else {
                tc::mbar_arrive_expect_tx(&bar->a_full, A_BYTES);
                for (int plane = 0; plane < 2; ++plane)
                    for (int kb = 0; kb < NKB; ++kb)
                        tc::tma_load_3d(sA + (plane * NKB + kb) * TILE_BYTES, plane ? &tm_m2l : &tm_m2h, &bar->a_full, kb * 64, m0, b);
            }
            uint32_t u = 0;
            for (int t = t0; t < t1; ++t)
                for (int kb = 0; kb < NKB; ++kb)
                    for (int plane = 0; plane < 2; ++plane, ++u) {
                        const uint32_t stage = u % UNITS, ph = (u / UNITS) & 1;
                        tc::mbar_wait(&bar->b_empty[stage], ph ^ 1);
                        if (CG == 2) {
                            if (rank == 0) tc::mbar_arrive_expect_tx(&bar->b_full[stage], 2 * UNIT_BYTES);
                            tc::tma_load_3d_cg2(sB + stage * UNIT_BYTES, plane ? &tm_m3l : &tm_m3h, tc::mapa(tc::smem_u32(&bar->b_full[stage]), 0),
                                                kb * 64, t * BN + (int)rank * Cfg::B_ROWS, b);
                        } else {
                            tc::mbar_arrive_expect_tx(&bar->b_full[stage], UNIT_BYTES);
                            tc::tma_load_3d(sB + stage * UNIT_BYTES, plane ? &tm_m3l : &tm_m3h, &bar->b_full[stage], kb * 64, t * BN, b);
                        }
                    }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ================= MMA issuer =================
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = tc::umma_idesc_f16(BM * CG, BN);
            auto mma = [](uint32_t dd, uint64_t a, uint64_t bq, uint32_t acc) { if (CG == 2) tc::umma_f16_cg2(dd, a, bq, idesc, acc); else tc::umma_f16(dd, a, bq, idesc, acc); };
            auto commit = [](uint64_t *barrier) { if (CG == 2) tc::umma_commit_cg2(barrier); else tc::umma_commit(barrier); };
            tc::mbar_wait(&bar->a_full, 0);
            tc::tcgen05_fence_after();
            const uint32_t a_base = tc::smem_u32(sA), b_base = tc::smem_u32(sB);
            uint32_t u = 0;
            int j = 0;
            for (int t = t0; t < t1; ++t, ++j) {
                const uint32_t as = j & 1, aph = (j >> 1) & 1;
                tc::mbar_wait(&bar->acc_empty[as], aph ^ 1);
                tc::tcgen05_fence_after();
                const uint32_t d = tmem_base + as * BN;
                if (F8) {
                    constexpr uint32_t id_lh = tc::umma_idesc_f8(BM * CG, BN, tc::kF8E5M2, tc::kF8E4M3);     // A lo (e5m2) x B hi (e4m3)
                    constexpr uint32_t id_hl = tc::umma_idesc_f8(BM * CG, BN, tc::kF8E4M3, tc::kF8E5M2);     // A hi (e4m3) x B lo (e5m2)
                    for (int q = 0; q < 8; ++q, ++u) {
                        const uint32_t stage = u % UNITS, ph = (u / UNITS) & 1;
                        tc::mbar_wait(&bar->b_full[stage], ph);
                        tc::tcgen05_fence_after();
                        const uint64_t bd = tc::umma_desc_k128(b_base + stage * UNIT_BYTES);
                        // A block q: hi fp16 k-block q (q < 4); lo e5m2 half q - 4 against B hi e4m3; hi e4m3 half q - 6 against B lo e5m2
                        const uint64_t ad = tc::umma_desc_k128(a_base + q * TILE_BYTES);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {            // 32 bytes per k-step: 16 fp16 or 32 bytes of fp8
                            if (p.dbg & 8) continue;
                            if (q < 4) tc::umma_f16_cg2(d, ad + 2 * k, bd + 2 * k, idesc, (q | k) != 0);
                            else tc::umma_f8_cg2(d, ad + 2 * k, bd + 2 * k, q < 6 ? id_lh : id_hl, 1);
                        }
                        commit(&bar->b_empty[stage]);
                    }
                }
                for (int kb = 0; kb < (F8 ? 0 : NKB); ++kb)
                    for (int plane = 0; plane < 2; ++plane, ++u) {
                        const uint32_t stage = u % UNITS, ph = (u / UNITS) & 1;
                        tc::mbar_wait(&bar->b_full[stage], ph);
                        tc::tcgen05_fence_after();
                        const uint64_t bd = tc::umma_desc_k128(b_base + stage * UNIT_BYTES);
                        const uint64_t ah = tc::umma_desc_k128(a_base + kb * TILE_BYTES);
                        const uint64_t al = tc::umma_desc_k128(a_base + (NKB + kb) * TILE_BYTES);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {            // UMMA_K = 16 fp16 = 32 bytes -> +2 in the address field
                            if (p.dbg & 8) continue;
                            if (plane == 0) {
                                mma(d, ah + 2 * k, bd + 2 * k, (kb | k) != 0);
                                mma(d, al + 2 * k, bd + 2 * k, 1);
                            } else {
                                mma(d, ah + 2 * k, bd + 2 * k, 1);
                            }
                        }
                        commit(&bar->b_empty[stage]);             // frees the ring unit (in both CTAs of a pair) when these MMAs retire
                    }
                commit(&bar->acc_full[as]);
            }
        }
        __syncwarp();
    } else {
        // ================= epilogue (warps 0-7): TMEM lane quadrant = warp % 4 (thread = query row), column half = warp / 4 ===
        const int quad = warp & 3, half = warp >> 2;
        const int row = m0 + quad * 32 + lane;
        const bool row_ok = row < p.N2;
        const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
        // PASS 2 row selection: running maximum of the raw conf bits (conf >= +0: unsigned order = float order) and the first column of
        // the 16-column block that holds it (strictly greater: the earliest block wins ties); k_sim_finalize_max finds the exact column
        // by re-reading those 16 scores from conf
        uint32_t row_best = 0u, row_blk = 0u;
        // PASS 2: conf = exp2(2*c1*acc - 2*c0 + log2(1/R_n) + log2(1/C_m)); rows / columns past the end carry -inf => conf = +0
        float rowc = 0.f;
        if (PASS == 2) rowc = row_ok ? p.lgR[(long long)b * p.N2 + row] - 2.f * p.c0 : -INFINITY;
        float *crow = PASS == 2 ? p.conf + ((long long)b * p.N2 + (row_ok ? row : 0)) * p.N3 : nullptr;
        const bool vec_ok = (p.N3 & 3) == 0;
        // PASS 2 column selection reads the staged 32 x 16 block back transposed: lane = (column c, row parity par), 16 rows each.
        // Keys: conf bits with the 7 low mantissa bits replaced by the inverted row index inside the CTA's 128 rows (ties within
        // 2^-16 relative resolve to the smaller row; the exact score is read back from conf by k_mutual_select)
        const int tc_c = lane & 15, tc_par = lane >> 4;
        const uint32_t rinv0 = 127u - (uint32_t)(quad * 32 + tc_par);       // row 2j + par of the warp's 32 rows: key index rinv0 - 2j
        uint32_t tc_off[4];                                                 // byte offset of (row 2j + par, column c) for j & 3 = 0..3, minus 128 j
#pragma unroll
        for (int k = 0; k < 4; ++k) tc_off[k] = (uint32_t)(tc_par * 64 + (((tc_c >> 2) ^ k) << 4) + (tc_c & 3) * 4);
        // fold of the four quadrants' column partials of tile (tt, buffer jj), thread = column, fixed order
        auto fold_columns = [&](int tt, int jj) {
            const uint32_t *scr = scratch + (jj & 1) * 4 * BN;
            const int tid = threadIdx.x, col = tt * BN + tid;
            if (tid < BN && col < p.N3 && tile_ok) {
                if (PASS == 1) {
                    const float *sf = reinterpret_cast<const float *>(scr);
                    p.colsum_part[((long long)b * p.tiles_m + row_tile) * p.N3 + col] = ((sf[tid] + sf[BN + tid]) + sf[2 * BN + tid]) + sf[3 * BN + tid];
                } else {
                    uint32_t k = scr[tid];
#pragma unroll
                    for (int w = 1; w < 4; ++w) k = scr[w * BN + tid] > k ? scr[w * BN + tid] : k;
                    p.colmax_part[((long long)b * p.tiles_m + row_tile) * p.N3 + col] =
                        ((unsigned long long)(k & 0xFFFFFF80u) << 32) | (unsigned long long)(0xffffffffu - (unsigned)(m0 + 127 - (int)(k & 127u)));
                }
            }
        };
        auto load_lgc = [&](int tt) -> float {           // column term of tile tt for this thread (threads 0..BN-1)
            const int col = tt * BN + (int)threadIdx.x;
            return (PASS == 2 && threadIdx.x < BN && tt < t1 && col < p.N3) ? p.lgC[(long long)b * p.N3 + col] : -INFINITY;
        };
        float lg_next = load_lgc(t0);
        int j = 0;
        for (int t = t0; t < t1; ++t, ++j) {
            const uint32_t as = j & 1, aph = (j >> 1) & 1;
            uint32_t *scr = scratch + (j & 1) * 4 * BN;
            float *lgc = s_lgc + (j & 1) * BN;
            if (PASS == 2 && threadIdx.x < BN) lgc[threadIdx.x] = lg_next;
            epi_bar_sync();                       // one barrier per tile: lgc(j) is visible and the partials of tile j-1 are complete
            if (j > 0) fold_columns(t - 1, j - 1);
            lg_next = load_lgc(t + 1);            // in flight during this tile
            tc::mbar_wait(&bar->acc_full[as], aph);
            tc::tcgen05_fence_after();
            float row_sum = 0.f;
#pragma unroll 1
            for (int g = 0; g < GROUPS; ++g) {
                // two 32-column chunks are fetched up front; after the last pair the accumulator goes back to the MMA warp
                float v[2][32];
                tc::tmem_ld_32x32(lane_base + as * BN + half * HALF + g * 64, v[0]);
                tc::tmem_ld_32x32(lane_base + as * BN + half * HALF + g * 64 + 32, v[1]);
                tc::tmem_ld_wait();
                if (g == GROUPS - 1) {
                    tc::tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) { if (CG == 2) tc::mbar_arrive_cluster(tc::mapa(tc::smem_u32(&bar->acc_empty[as]), 0)); else tc::mbar_arrive(&bar->acc_empty[as]); }
                }
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int hcol = g * 64 + cc * 32;             // column inside this warp's half
                    const int ccol = half * HALF + hcol;           // column inside the tile
                    const int col0 = t * BN + ccol;
                    float *vv = v[cc];
                    if (PASS == 1) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const float e = tc::ex2(fmaf(vv[i], p.c1, -p.c0));
                            vv[i] = (row_ok && col0 + i < p.N3) ? e : 0.f;
                        }
#pragma unroll
                        for (int i = 0; i < 32; ++i) row_sum += vv[i];
                        const float cs = transpose_reduce_sum(vv, lane);
                        reinterpret_cast<float *>(scr)[quad * BN + ccol + lane] = cs;
                    } else {
                        uint8_t *stg = s_stg + warp * STG_WARP_BYTES;
#pragma unroll
                        for (int hc = 0; hc < 2; ++hc) {
                            // 16 columns of arithmetic, then their store: the TMA engine reads the staging buffer of the previous
                            // 16 columns while these are computed
                            // (no experiment switch inside this loop: a branch per score serialises FFMA -> MUFU -> ...)
                            float l[16];
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const float4 l4 = *reinterpret_cast<const float4 *>(lgc + ccol + 16 * hc + 4 * q);
                                l[4 * q] = l4.x; l[4 * q + 1] = l4.y; l[4 * q + 2] = l4.z; l[4 * q + 3] = l4.w;
                            }
                            uint32_t bm = 0u;
#pragma unroll
                            for (int e = 0; e < 16; ++e) {
                                const int i = 16 * hc + e;
                                vv[i] = tc::ex2(fmaf(vv[i], p.c2, rowc + l[e]));
                                bm = max(bm, __float_as_uint(vv[i]));
                            }
                            if (bm > row_best) { row_best = bm; row_blk = (uint32_t)(col0 + 16 * hc); }
                            // stage 32 rows x 16 columns in (64-byte swizzled) shared memory: one coalesced TMA store per block (rows /
                            // columns past the end of conf are clipped by the TMA unit), and the source of the transposed read below
                            if (lane == 0) tc::tma_store_wait_read0();        // the previous store has read the staging buffer
                            __syncwarp();
#pragma unroll
                            for (int q = 0; q < 4; ++q)
                                *reinterpret_cast<float4 *>(stg + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4)) =
                                    make_float4(vv[hc * 16 + 4 * q], vv[hc * 16 + 4 * q + 1], vv[hc * 16 + 4 * q + 2], vv[hc * 16 + 4 * q + 3]);
                            tc::fence_proxy_async();
                            __syncwarp();
                            if (p.tma_store && lane == 0 && !(p.dbg & 1)) {
                                tc::tma_store_3d(&tm_conf, stg, col0 + hc * 16, m0 + quad * 32, b);
                                tc::tma_store_commit();
                            }
                            // column keys: 16 rows of one parity per lane (even rows sit in banks 0-15, odd rows in 16-31: conflict-free)
                            uint32_t ck = 0u;
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const uint32_t w = *reinterpret_cast<const uint32_t *>(stg + tc_off[j & 3] + 128 * j);
                                ck = max(ck, (w & 0xFFFFFF80u) | (rinv0 - 2u * j));
                            }
                            ck = max(ck, __shfl_xor_sync(0xffffffffu, ck, 16));
                            if (lane < 16) scr[quad * BN + ccol + 16 * hc + lane] = ck;
                        }
                        if (!p.tma_store && row_ok) {
                            if (vec_ok && col0 + 31 < p.N3) {
#pragma unroll
                                for (int q = 0; q < 8; ++q)
                                    *reinterpret_cast<float4 *>(crow + col0 + 4 * q) = make_float4(vv[4 * q], vv[4 * q + 1], vv[4 * q + 2], vv[4 * q + 3]);
                            } else {
                                for (int i = 0; i < 32; ++i) if (col0 + i < p.N3) crow[col0 + i] = vv[i];
                            }
                        }
                    }
                }
            }
            if (PASS == 1) {
                // per-(tile, column half) partial: the fold over partials is split-invariant
                if (row_ok) p.rowsum_part[((long long)b * (p.tiles_n * 2) + t * 2 + half) * p.N2 + row] = row_sum;
            }
        }
        epi_bar_sync();
        fold_columns(t1 - 1, j - 1);
        if (PASS == 2 && lane == 0) tc::tma_store_wait_all();
        if (PASS == 2 && row_ok)      // (value bits, inverted first column of the winning 16-column block)
            p.rowmax_part[((long long)b * (gridDim.y * 2) + split * 2 + half) * p.N2 + row] =
                ((unsigned long long)row_best << 32) | (unsigned long long)(0xffffffffu - row_blk);
    }
    tc::tcgen05_fence_before();
    if (CG == 2) tc::cluster_sync(); else __syncthreads();                // nothing is in flight to or from the peer any more
    if (warp == 9) { if (CG == 2) tc::tmem_dealloc_cg2<Cfg::TMEM_COLS>(tmem_base); else tc::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base); }
}

// Operand planes of a row-major fp32 tensor, x scaled by `scale`: hi = fp16(x), lo = fp16(x - hi).
// F8 = 0: hi and lo as fp16 arrays.  F8 = 1: hi as fp16, lo as e5m2 bytes (`lo`), hi as e4m3 bytes (`hx`) for the fp8 cross terms.
template <int F8>
__device__ __forceinline__ void store_planes(float xs, long long i, __half *hi, void *lo, uint8_t *hx) {
    __half h, l;
    tc::split_f16(xs, h, l);
    hi[i] = h;
    if (F8) {
        reinterpret_cast<uint8_t *>(lo)[i] = (uint8_t)__nv_cvt_float_to_fp8(__half2float(l), __NV_SATFINITE, __NV_E5M2);
        hx[i] = (uint8_t)__nv_cvt_float_to_fp8(__half2float(h), __NV_SATFINITE, __NV_E4M3);
    } else {
        reinterpret_cast<__half *>(lo)[i] = l;
    }
}

template <int F8>
__global__ void k_split_planes(const float *__restrict__ x, __half *__restrict__ hi, void *__restrict__ lo, uint8_t *__restrict__ hx, long long n, float scale) {
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const float4 v = *reinterpret_cast<const float4 *>(x + i);
    store_planes<F8>(v.x * scale, i, hi, lo, hx); store_planes<F8>(v.y * scale, i + 1, hi, lo, hx);
    store_planes<F8>(v.z * scale, i + 2, hi, lo, hx); store_planes<F8>(v.w * scale, i + 3, hi, lo, hx);
}

// F.normalize (x / max(|x|, 1e-12), GATs_SuperGlue.py:245-246) fused into the operand split: one warp per 256-wide row, the same
// summation order and quotient as k_l2_normalize_rows, so the planes are bit-identical to normalising first and splitting after
template <int F8>
__global__ void k_normalize_split_planes(const float *__restrict__ x, __half *__restrict__ hi, void *__restrict__ lo, uint8_t *__restrict__ hx,
                                         long long rows, float scale) {
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const float *src = x + row * 256;
    float v[8], ss = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { v[k] = src[lane + 32 * k]; ss += v[k] * v[k]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float den = fmaxf(sqrtf(ss), 1e-12f);
#pragma unroll
    for (int k = 0; k < 8; ++k) store_planes<F8>((v[k] / den) * scale, row * 256 + lane + 32 * k, hi, lo, hx);
}

}  // namespace simtc
}  // namespace opb
