// Similarity + dual-softmax passes with SIXTEEN epilogue warps (included by sim_dual_softmax.cu): the production form.
//
// Same tiling and MMA schedule as k_sim_tc<PASS, 2, F8> (sim_tc.cuh: CTA pair, M256 x N256 tiles, A resident, B through a ring of four
// 16 KB units, double-buffered 2 x 256-column TMEM accumulators), but the epilogue is the part that bounds both passes (ncu of the
// eight-warp form: issue slots 35 % busy in pass 2 with two epilogue warps per scheduler, every warp a chain of
// tcgen05.ld -> FFMA -> MUFU -> STS -> fence -> TMA store -> LDS), so here
//   * warps 0-15 are epilogue warps: TMEM lane quadrant = warp % 4 (thread = query row), column QUARTER = warp / 4 (64 columns of the
//     tile, two 32-column chunks) -- four warps per scheduler to overlap each other's dependency chains;
//   * no block-wide barrier and no shared scratch in the tile loop: every warp writes its own partials to global memory (column sums /
//     column keys per (row tile, quadrant), row sums / row keys per (column tile | split, quarter)), the small finalize kernels fold
//     them in a fixed order -- bit-reproducible, and a frame's bits do not depend on the batch it is in;
//   * pass 2 takes the column terms log2(1/C_m) from two registers per lane (columns lane, lane + 32 of the warp's quarter, loaded one
//     tile ahead) through warp shuffles instead of a staged shared array.
// warp 16 = TMA producer, warp 17 = TMEM allocation + MMA issue (leader CTA only).
#pragma once
#include "sim_tc.cuh"

namespace opb {
namespace simtc16 {

using simtc::A_BYTES;
using simtc::BM;
using simtc::NKB;
using simtc::RING_BYTES;
using simtc::TILE_BYTES;

constexpr int EW = 16;                                // epilogue warps
constexpr int THREADS = (EW + 2) * 32;                // 576
constexpr int BN = 256, MAX_UNITS = 4, UNIT_BYTES = 128 * 128, B_ROWS = 128, QCOLS = 64;
constexpr int STG_BUF_BYTES = 32 * 16 * 4;            // one staging buffer: 32 rows x 16 fp32 columns, or 16 rows x 32 columns (pass 2)
// NU = ring units of 16 KB (B operand), DB = 1: two staging buffers per epilogue warp.  Shared memory is full with A resident (128 KB):
// 4 units + one buffer per warp (64 + 32 KB) or 2 units + two buffers per warp (32 + 64 KB).
template <int PASS, int NU = 4, int DB = 0>
constexpr int smem_bytes() { return A_BYTES + NU * UNIT_BYTES + (PASS == 2 ? EW * STG_BUF_BYTES * (DB ? 2 : 1) : 0) + 256 + 1024; }   // + barriers + alignment slack

struct Params {
    int N2, N3, tiles_n, tiles_m, tiles_per_split;
    float c1, c0, c2;                                 // pass 1: e = ex2(acc c1 - c0); pass 2: conf = ex2(acc c2 - 2 c0 + lgR + lgC)
    float *rowsum_part;                               // [B][tiles_n * 4][N2]
    float *colsum_part;                               // [B][tiles_m * 4][N3]
    const float *lgR, *lgC;                           // log2(1/R_n), log2(1/C_m)
    unsigned long long *rowmax_part;                  // [B][splits * 4][N2]: (conf bits, inverted first column of the winning 16-column block)
    uint32_t *colkey_part;                            // [B][tiles_m * 4][N3]: conf bits, low 7 bits = 127 - row inside the CTA's 128 rows
    float *conf;
    int tma_store;
    const int *n2_counts;                             // ragged batches: [B] valid query rows per frame (rows past the count are padding: conf rows of
                                                      // zeros, no part in the column sums / column maxima); null = N2 rows everywhere
    int dbg;                                          // OPB_SIM_DBG (timing experiments): 1 = no conf stores, 4 = stores into one L2-resident patch, 8 = no MMAs
};

struct Barriers {
    uint64_t a_full, b_full[MAX_UNITS], b_empty[MAX_UNITS], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

template <int PASS, int F8, int WIDE = 0, int NU = 4, int DB = 0>
__global__ void __launch_bounds__(THREADS, 1)
k_sim16(const __grid_constant__ CUtensorMap tm_m2h, const __grid_constant__ CUtensorMap tm_m2l, const __grid_constant__ CUtensorMap tm_m2x,
        const __grid_constant__ CUtensorMap tm_m3h, const __grid_constant__ CUtensorMap tm_m3l, const __grid_constant__ CUtensorMap tm_m3x,
        const __grid_constant__ CUtensorMap tm_conf, const Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *sA = smem, *sB = smem + A_BYTES;
    constexpr int UNITS = NU, STG_WARP_BYTES = STG_BUF_BYTES * (DB ? 2 : 1);
    uint8_t *s_stg = smem + A_BYTES + NU * UNIT_BYTES;
    Barriers *bar = reinterpret_cast<Barriers *>(smem + A_BYTES + NU * UNIT_BYTES + (PASS == 2 ? EW * STG_WARP_BYTES : 0));

    const uint32_t rank = tc::cluster_ctarank();                          // 0 = leader of the pair (row tiles 2i, 2i + 1)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int split = blockIdx.y, row_tile = blockIdx.x, b = blockIdx.z;
    const bool tile_ok = row_tile < p.tiles_m;                            // false: ghost row tile completing a pair
    const int m0 = row_tile * BM;
    const int t0 = split * p.tiles_per_split;
    const int t1 = min(p.tiles_n, t0 + p.tiles_per_split);
    if (t0 >= t1) return;                                                 // uniform for the whole pair

    if (threadIdx.x == 0) {
        tc::mbar_init(&bar->a_full, 1);
        for (int i = 0; i < UNITS; ++i) { tc::mbar_init(&bar->b_full[i], 1); tc::mbar_init(&bar->b_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&bar->acc_full[i], 1); tc::mbar_init(&bar->acc_empty[i], EW * 2); }
        tc::fence_barrier_init();
    }
    if (warp == EW + 1) tc::tmem_alloc_cg2<512>(&bar->tmem_base);
    if (warp == EW && lane == 0) {
        tc::tma_prefetch_desc(&tm_m2h); tc::tma_prefetch_desc(&tm_m2l); tc::tma_prefetch_desc(&tm_m3h); tc::tma_prefetch_desc(&tm_m3l);
        if (F8) { tc::tma_prefetch_desc(&tm_m2x); tc::tma_prefetch_desc(&tm_m3x); }
        if (PASS == 2 && p.tma_store) tc::tma_prefetch_desc(&tm_conf);
    }
    tc::tcgen05_fence_before();
    tc::cluster_sync();                                                   // the peer's barriers exist before anything is sent to them
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = bar->tmem_base;

    if (warp == EW) {
        // ================= TMA producer (as k_sim_tc) =================
        if (lane == 0) {
            const uint32_t af = tc::mapa(tc::smem_u32(&bar->a_full), 0);
            if (rank == 0) tc::mbar_arrive_expect_tx(&bar->a_full, 2 * A_BYTES);
            if (F8) {
                // A resident: hi fp16 (4 k-blocks of 64), then lo e5m2 and hi e4m3 (2 k-blocks of 128 bytes each)
                for (int kb = 0; kb < NKB; ++kb) tc::tma_load_3d_cg2(sA + kb * TILE_BYTES, &tm_m2h, af, kb * 64, m0, b);
                for (int j = 0; j < 2; ++j) tc::tma_load_3d_cg2(sA + (4 + j) * TILE_BYTES, &tm_m2l, af, j * 128, m0, b);
                for (int j = 0; j < 2; ++j) tc::tma_load_3d_cg2(sA + (6 + j) * TILE_BYTES, &tm_m2x, af, j * 128, m0, b);
            } else {
                for (int plane = 0; plane < 2; ++plane)
                    for (int kb = 0; kb < NKB; ++kb)
                        tc::tma_load_3d_cg2(sA + (plane * NKB + kb) * TILE_BYTES, plane ? &tm_m2l : &tm_m2h, af, kb * 64, m0, b);
            }
            uint32_t u = 0;
            for (int t = t0; t < t1; ++t)
                for (int q = 0; q < 8; ++q, ++u) {
                    const uint32_t stage = u % UNITS, ph = (u / UNITS) & 1;
                    tc::mbar_wait(&bar->b_empty[stage], ph ^ 1);
                    if (rank == 0) tc::mbar_arrive_expect_tx(&bar->b_full[stage], 2 * UNIT_BYTES);
                    // F8: 4 x hi fp16, 2 x hi e4m3, 2 x lo e5m2.  fp16 form: (k-block, plane) pairs, plane fastest
                    const CUtensorMap *tm = F8 ? (q < 4 ? &tm_m3h : (q < 6 ? &tm_m3x : &tm_m3l)) : ((q & 1) ? &tm_m3l : &tm_m3h);
                    const int c0 = F8 ? (q < 4 ? q * 64 : (q & 1) * 128) : (q >> 1) * 64;
                    tc::tma_load_3d_cg2(sB + stage * UNIT_BYTES, tm, tc::mapa(tc::smem_u32(&bar->b_full[stage]), 0), c0, t * BN + (int)rank * B_ROWS, b);
                }
        }
        __syncwarp();
    } else if (warp == EW + 1) {
        // ================= MMA issuer (as k_sim_tc) =================
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = tc::umma_idesc_f16(BM * 2, BN);
            constexpr uint32_t id_lh = tc::umma_idesc_f8(BM * 2, BN, tc::kF8E5M2, tc::kF8E4M3);     // A lo (e5m2) x B hi (e4m3)
            constexpr uint32_t id_hl = tc::umma_idesc_f8(BM * 2, BN, tc::kF8E4M3, tc::kF8E5M2);     // A hi (e4m3) x B lo (e5m2)
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
                for (int q = 0; q < 8; ++q, ++u) {
                    const uint32_t stage = u % UNITS, ph = (u / UNITS) & 1;
                    tc::mbar_wait(&bar->b_full[stage], ph);
                    tc::tcgen05_fence_after();
                    const uint64_t bd = tc::umma_desc_k128(b_base + stage * UNIT_BYTES);
                    if (F8) {
                        // A block q: hi fp16 k-block q (q < 4); lo e5m2 half q - 4 against B hi e4m3; hi e4m3 half q - 6 against B lo e5m2
                        const uint64_t ad = tc::umma_desc_k128(a_base + q * TILE_BYTES);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {            // 32 bytes per k-step: 16 fp16 or 32 bytes of fp8
                            if (p.dbg & 8) continue;
                            if (q < 4) tc::umma_f16_cg2(d, ad + 2 * k, bd + 2 * k, idesc, (q | k) != 0);
                            else tc::umma_f8_cg2(d, ad + 2 * k, bd + 2 * k, q < 6 ? id_lh : id_hl, 1);
                        }
                    } else {
                        const int kb = q >> 1, plane = q & 1;
                        const uint64_t ah = tc::umma_desc_k128(a_base + kb * TILE_BYTES);
                        const uint64_t al = tc::umma_desc_k128(a_base + (NKB + kb) * TILE_BYTES);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            if (p.dbg & 8) continue;
                            if (plane == 0) {
                                tc::umma_f16_cg2(d, ah + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
                                tc::umma_f16_cg2(d, al + 2 * k, bd + 2 * k, idesc, 1);
                            } else {
                                tc::umma_f16_cg2(d, ah + 2 * k, bd + 2 * k, idesc, 1);
                            }
                        }
                    }
                    tc::umma_commit_cg2(&bar->b_empty[stage]);        // frees the ring unit in both CTAs when these MMAs retire
                }
                tc::umma_commit_cg2(&bar->acc_full[as]);
            }
        }
        __syncwarp();
    } else {
        // ================= epilogue (warps 0-15): TMEM lane quadrant = warp % 4 (thread = query row), column quarter = warp / 4 =========
        const int quad = warp & 3, qtr = warp >> 2;
        const int row = m0 + quad * 32 + lane;
        const int n2 = p.n2_counts ? min(p.N2, __ldg(p.n2_counts + b)) : p.N2;       // valid query rows of this frame
        const bool row_ok = row < n2, row_in = row < p.N2;
        const bool rows_full = m0 + BM <= n2;                             // warp-uniform
        const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(qtr * QCOLS);
        const long long part_row = ((long long)b * p.tiles_m + row_tile) * 4 + quad;     // this warp's row of the per-(row tile, quadrant) partials
        const uint32_t acc_empty0 = tc::mapa(tc::smem_u32(&bar->acc_empty[0]), 0), acc_empty1 = tc::mapa(tc::smem_u32(&bar->acc_empty[1]), 0);
        auto release_acc = [&](uint32_t as) {
            tc::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive_cluster(as ? acc_empty1 : acc_empty0);
        };
        if (PASS == 1) {
            int j = 0;
            for (int t = t0; t < t1; ++t, ++j) {
                const uint32_t as = j & 1, aph = (j >> 1) & 1;
                const int colq = t * BN + qtr * QCOLS;                     // first column of this warp's quarter
                const bool full = rows_full && colq + QCOLS <= p.N3;      // no element of the quarter needs masking
                tc::mbar_wait(&bar->acc_full[as], aph);
                tc::tcgen05_fence_after();
                float row_sum = 0.f;
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    float v[32];
                    tc::tmem_ld_32x32(lane_base + as * BN + cc * 32, v);
                    tc::tmem_ld_wait();
                    if (cc == 1) release_acc(as);
                    const int col0 = colq + cc * 32;
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = tc::ex2(fmaf(v[i], p.c1, -p.c0));
                    if (!full) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = (row_ok && col0 + i < p.N3) ? v[i] : 0.f;
                    }
#pragma unroll
                    for (int i = 0; i < 32; ++i) row_sum += v[i];
                    const float cs = simtc::transpose_reduce_sum(v, lane);
                    if (tile_ok && col0 + lane < p.N3) p.colsum_part[part_row * p.N3 + col0 + lane] = cs;
                }
                // per-(tile, quarter) partial: the fold over partials does not depend on how the tiles are split over CTAs
                if (row_ok) p.rowsum_part[((long long)b * (p.tiles_n * 4) + t * 4 + qtr) * p.N2 + row] = row_sum;
            }
        } else {
            // conf = exp2(2 c1 acc - 2 c0 + log2(1/R_n) + log2(1/C_m)); rows / columns past the end carry -inf => conf = +0
            const float rowc = row_ok ? p.lgR[(long long)b * p.N2 + row] - 2.f * p.c0 : -INFINITY;
            float *crow = p.conf + ((long long)b * p.N2 + (row_in ? row : 0)) * p.N3;
            const bool vec_ok = (p.N3 & 3) == 0;
            // row selection: running maximum of the raw conf bits (conf >= +0: unsigned order = float order) and the first column of the
            // 16-column block that holds it (strictly greater: the earliest block wins ties); k_sim_finalize_max finds the exact column
            uint32_t row_best = 0u, row_blk = 0u;
            // column selection reads the staged 32 x 16 block back transposed: lane = (column c, row parity par), 16 rows each.  Keys: conf
            // bits with the 7 low mantissa bits replaced by the inverted row index inside the CTA's 128 rows (ties within 2^-16 relative
            // resolve to the smaller row; the exact score is read back from conf by k_mutual_select)
            const int tc_c = lane & 15, tc_par = lane >> 4;
            const uint32_t rinv0 = 127u - (uint32_t)(quad * 32 + tc_par);
            uint32_t tc_off[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) tc_off[k] = (uint32_t)(tc_par * 64 + (((tc_c >> 2) ^ k) << 4) + (tc_c & 3) * 4);
            uint8_t *stg0 = s_stg + warp * STG_WARP_BYTES;
            uint8_t *stg = stg0;                                           // (the narrow form uses one buffer)
            auto load_lgc = [&](int tt, int k) -> float {                 // column term of column (lane + 32 k) of this warp's quarter of tile tt
                const int col = tt * BN + qtr * QCOLS + 32 * k + lane;
                return (tt < t1 && col < p.N3) ? __ldg(p.lgC + (long long)b * p.N3 + col) : -INFINITY;
            };
            float lg_next0 = load_lgc(t0, 0), lg_next1 = load_lgc(t0, 1);
            int j = 0;
            for (int t = t0; t < t1; ++t, ++j) {
                const uint32_t as = j & 1, aph = (j >> 1) & 1;
                const float lg0 = lg_next0, lg1 = lg_next1;
                lg_next0 = load_lgc(t + 1, 0); lg_next1 = load_lgc(t + 1, 1);     // in flight during this tile
                const int colq = t * BN + qtr * QCOLS;
                tc::mbar_wait(&bar->acc_full[as], aph);
                tc::tcgen05_fence_after();
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    float v[32];
                    tc::tmem_ld_32x32(lane_base + as * BN + cc * 32, v);
                    tc::tmem_ld_wait();
                    if (cc == 1) release_acc(as);
                    const int col0 = colq + cc * 32;
                    const float lgc = cc ? lg1 : lg0;
                    if (WIDE) {
                        // 32 columns of arithmetic; then the chunk leaves as two TMA stores of 16 rows x 128 BYTES (whole 128-byte lines of
                        // conf per row instead of 64-byte halves), staged by the half-warp that owns the rows; the column keys are read
                        // back transposed from the same staging buffer (lane = column, 16 rows per half, bank-conflict-free under the
                        // 128-byte swizzle)
                        uint32_t bm0 = 0u, bm1 = 0u;
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            v[i] = tc::ex2(fmaf(v[i], p.c2, rowc + __shfl_sync(0xffffffffu, lgc, i)));
                            if (i < 16) bm0 = max(bm0, __float_as_uint(v[i])); else bm1 = max(bm1, __float_as_uint(v[i]));
                        }
                        if (bm0 > row_best) { row_best = bm0; row_blk = (uint32_t)col0; }
                        if (bm1 > row_best) { row_best = bm1; row_blk = (uint32_t)(col0 + 16); }
                        uint32_t ck = 0u;
#pragma unroll
                        for (int hf = 0; hf < 2; ++hf) {
                            // DB: the two halves of a chunk alternate between the warp's two buffers, so a half only waits for the
                            // store issued one whole chunk earlier (at most one younger store still reading); one buffer: for the previous half's
                            uint8_t *stg = stg0 + (DB ? hf * STG_BUF_BYTES : 0);
                            if (lane == 0) { if (DB) tc::tma_store_wait_read1(); else tc::tma_store_wait_read0(); }
                            __syncwarp();
                            if ((lane >> 4) == hf) {
#pragma unroll
                                for (int q = 0; q < 8; ++q)
                                    *reinterpret_cast<float4 *>(stg + (lane & 15) * 128 + ((q ^ (lane & 7)) << 4)) =
                                        make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                            }
                            tc::fence_proxy_async();
                            __syncwarp();
                            if (p.tma_store && lane == 0 && !(p.dbg & 1)) {
                                if (p.dbg & 4) tc::tma_store_3d(&tm_conf, stg, col0 & 255, quad * 32 + hf * 16, 0);   // timing experiment: every store lands in one L2-resident patch
                                else tc::tma_store_3d(&tm_conf, stg, col0, m0 + quad * 32 + hf * 16, b);
                                tc::tma_store_commit();
                            }
                            const uint32_t rinv = 127u - (uint32_t)(quad * 32 + hf * 16);
#pragma unroll
                            for (int r = 0; r < 16; ++r) {
                                const uint32_t w = *reinterpret_cast<const uint32_t *>(stg + r * 128 + ((((lane >> 2) ^ (r & 7)) << 4) | ((lane & 3) << 2)));
                                ck = max(ck, (w & 0xFFFFFF80u) | (rinv - (uint32_t)r));
                            }
                        }
                        if (tile_ok && col0 + lane < p.N3) p.colkey_part[part_row * p.N3 + col0 + lane] = ck;
                    } else {
#pragma unroll
                    for (int hc = 0; hc < 2; ++hc) {
                        // 16 columns of arithmetic, then their store: the TMA engine reads the staging buffer while the column keys are built
                        uint32_t bm = 0u;
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            const int i = 16 * hc + e;
                            v[i] = tc::ex2(fmaf(v[i], p.c2, rowc + __shfl_sync(0xffffffffu, lgc, i)));
                            bm = max(bm, __float_as_uint(v[i]));
                        }
                        if (bm > row_best) { row_best = bm; row_blk = (uint32_t)(col0 + 16 * hc); }
                        // stage 32 rows x 16 columns in (64-byte swizzled) shared memory: one coalesced TMA store per block (rows / columns
                        // past the end of conf are clipped by the TMA unit), and the source of the transposed read below
                        if (lane == 0) tc::tma_store_wait_read0();        // the previous store has read the staging buffer
                        __syncwarp();
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            *reinterpret_cast<float4 *>(stg + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4)) =
                                make_float4(v[hc * 16 + 4 * q], v[hc * 16 + 4 * q + 1], v[hc * 16 + 4 * q + 2], v[hc * 16 + 4 * q + 3]);
                        tc::fence_proxy_async();
                        __syncwarp();
                        if (p.tma_store && lane == 0 && !(p.dbg & 1)) {
                            tc::tma_store_3d(&tm_conf, stg, col0 + hc * 16, m0 + quad * 32, b);
                            tc::tma_store_commit();
                        }
                        // column keys: 16 rows of one parity per lane (even rows sit in banks 0-15, odd rows in 16-31: conflict-free)
                        uint32_t ck = 0u;
#pragma unroll
                        for (int r = 0; r < 16; ++r) {
                            const uint32_t w = *reinterpret_cast<const uint32_t *>(stg + tc_off[r & 3] + 128 * r);
                            ck = max(ck, (w & 0xFFFFFF80u) | (rinv0 - 2u * r));
                        }
                        ck = max(ck, __shfl_xor_sync(0xffffffffu, ck, 16));
                        const int kcol = col0 + 16 * hc + lane;
                        if (lane < 16 && tile_ok && kcol < p.N3) p.colkey_part[part_row * p.N3 + kcol] = ck;
                    }
                    }
                    if (!p.tma_store && row_in) {
                        if (vec_ok && col0 + 31 < p.N3) {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                *reinterpret_cast<float4 *>(crow + col0 + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                        } else {
                            for (int i = 0; i < 32; ++i) if (col0 + i < p.N3) crow[col0 + i] = v[i];
                        }
                    }
                }
            }
            if (lane == 0) tc::tma_store_wait_all();
            if (row_in)      // (value bits, inverted first column of the winning 16-column block); padding rows of a ragged batch: (0, column 0)
                p.rowmax_part[((long long)b * (gridDim.y * 4) + split * 4 + qtr) * p.N2 + row] =
                    ((unsigned long long)row_best << 32) | (unsigned long long)(0xffffffffu - row_blk);
        }
    }
    tc::tcgen05_fence_before();
    tc::cluster_sync();                                                   // nothing is in flight to or from the peer any more
    if (warp == EW + 1) tc::tmem_dealloc_cg2<512>(tmem_base);
}

// column arg-maxima from the per-(row tile, quadrant) keys: partials in ascending row order, strictly greater on the masked conf bits,
// so ties (within 2^-16 relative) resolve to the smallest row -- first-maximum semantics, as torch.max(dim)
__global__ void k_sim_finalize_colkeys(const uint32_t *__restrict__ part, int parts, int n, int n_rows, int *__restrict__ idx, float *__restrict__ val) {
    const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t best = 0u;
    int row = 0;
    for (int t = 0; t < parts; ++t) {
        const uint32_t k = part[((long long)b * parts + t) * n + i];
        if ((k & 0xFFFFFF80u) > best) { best = k & 0xFFFFFF80u; row = (t >> 2) * BM + 127 - (int)(k & 127u); }
    }
    idx[(long long)b * n + i] = row < n_rows ? row : n_rows - 1;
    val[(long long)b * n + i] = __uint_as_float(best);
}

}  // namespace simtc16
}  // namespace opb
