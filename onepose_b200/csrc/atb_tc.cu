// (2) Source side of linear attention on the tensor cores -- reference linear_attention, GATs_SuperGlue.py:88-99.
//
// KV_h = sum_m K'_h[m]^T V_h[m] / M with V = X Wv^T + bv is re-associated as
//     KV_h = ( (K'_h^T X) Wv_h^T + Ksum_h bv_h^T ) / M ,
// so the only reduction over the M source points is G = K'^T X, a plain [256 x M] x [M x 256] GEMM (the V projection
// is never materialised), and the attention read-out + merge convolution (:98, :119) collapse into one per-frame matrix
//     P_b[o, (h,d)] = sum_q KV_h[d,q] Wm[o, (h,q)]
// applied to Q'' = Q' M / (Q'_h . Ksum_h + 1e-6) by the ordinary projection kernel (linear_tc.cuh, ZSCALE + WBATCHED).
//
//  k_atb_tc   : G partial of one chunk of points.  Both operands are row-major [points, 256] split-precision planes,
//               i.e. MN-major UMMA operands (K = point): TMA boxes of 32 points x 64 channels (SWIZZLE_128B), three
//               stages of 64 KB, tcgen05.mma M128 N256 K16 x 3 passes into two 256-column TMEM accumulators.
//  k_kv_post  : per (frame, head): fold the chunk partials and the Ksum partials in a fixed order, KV', P_b -> planes.
#include "common.cuh"
#include "tc_common.cuh"

namespace opb {
namespace atb {

constexpr int KB = 32;                         // points per k-block
constexpr int BLK = KB * 128;                  // one 64-channel block of one plane: 32 rows x 128 B = 4096
constexpr int PLANE = 4 * BLK;                 // 256 channels
constexpr int STAGE_BYTES = 4 * PLANE;         // A hi, A lo, B hi, B lo = 65536
constexpr int STAGES = 3;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 256 + 1024;
constexpr int THREADS = 320;
constexpr int kChunk = 1024;                   // points per CTA

struct Barriers { uint64_t full[STAGES], empty[STAGES], acc_full; uint32_t tmem_base; };

// MN-major SWIZZLE_128B operand: 64-element (128 B) atoms along M/N spaced LBO apart, 8-row groups along K spaced SBO apart
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}

__global__ void __launch_bounds__(THREADS, 1)
k_atb_tc(const __grid_constant__ CUtensorMap tm_ah, const __grid_constant__ CUtensorMap tm_al,
         const __grid_constant__ CUtensorMap tm_bh, const __grid_constant__ CUtensorMap tm_bl, int M, int nchunks,
         float *__restrict__ partial /* [B][nchunks][256][256] */) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    Barriers *bar = reinterpret_cast<Barriers *>(smem + STAGES * STAGE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = blockIdx.x, b = blockIdx.y;
    const int p0 = chunk * kChunk, p1 = min(M, p0 + kChunk);
    const int nkb = (p1 - p0 + KB - 1) / KB;

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; ++i) { tc::mbar_init(&bar->full[i], 1); tc::mbar_init(&bar->empty[i], 1); }
        tc::mbar_init(&bar->acc_full, 1);
        tc::fence_barrier_init();
    }
    if (warp == 9) tc::tmem_alloc<512>(&bar->tmem_base);
    if (warp == 8 && lane == 0) { tc::tma_prefetch_desc(&tm_ah); tc::tma_prefetch_desc(&tm_al); tc::tma_prefetch_desc(&tm_bh); tc::tma_prefetch_desc(&tm_bl); }
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = bar->tmem_base;

    if (warp == 8) {
        if (lane == 0) {
            for (int kb = 0; kb < nkb; ++kb) {
                const uint32_t stage = kb % STAGES, ph = (kb / STAGES) & 1;
                uint8_t *st = smem + stage * STAGE_BYTES;
                tc::mbar_wait(&bar->empty[stage], ph ^ 1);
                tc::mbar_arrive_expect_tx(&bar->full[stage], STAGE_BYTES);
                const int row = p0 + kb * KB;                 // rows past M are zero-filled by TMA
#pragma unroll
                for (int blk = 0; blk < 4; ++blk) {
                    tc::tma_load_3d(st + 0 * PLANE + blk * BLK, &tm_ah, &bar->full[stage], blk * 64, row, b);
                    tc::tma_load_3d(st + 1 * PLANE + blk * BLK, &tm_al, &bar->full[stage], blk * 64, row, b);
                    tc::tma_load_3d(st + 2 * PLANE + blk * BLK, &tm_bh, &bar->full[stage], blk * 64, row, b);
                    tc::tma_load_3d(st + 3 * PLANE + blk * BLK, &tm_bl, &bar->full[stage], blk * 64, row, b);
                }
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        if (lane == 0) {
            // D = f32, A = B = f16, both MN-major (bits 15, 16), M = 128, N = 256
            constexpr uint32_t idesc = tc::umma_idesc_f16(128, 256) | (1u << 15) | (1u << 16);
            const uint32_t base = tc::smem_u32(smem);
            for (int kb = 0; kb < nkb; ++kb) {
                const uint32_t stage = kb % STAGES, ph = (kb / STAGES) & 1;
                const uint32_t sb = base + stage * STAGE_BYTES;
                tc::mbar_wait(&bar->full[stage], ph);
                tc::tcgen05_fence_after();
#pragma unroll
                for (int ks = 0; ks < KB / 16; ++ks) {         // 16 points = two 8-row groups = 2048 B
                    const uint64_t bh = umma_desc_mn128(sb + 2 * PLANE + ks * 2048, BLK, 1024);
                    const uint64_t bl = umma_desc_mn128(sb + 3 * PLANE + ks * 2048, BLK, 1024);
#pragma unroll
                    for (int mh = 0; mh < 2; ++mh) {           // output rows (K' channels) mh*128 .. +127 = channel blocks 2mh, 2mh+1
                        const uint64_t ah = umma_desc_mn128(sb + 0 * PLANE + mh * 2 * BLK + ks * 2048, BLK, 1024);
                        const uint64_t al = umma_desc_mn128(sb + 1 * PLANE + mh * 2 * BLK + ks * 2048, BLK, 1024);
                        const uint32_t d = tmem_base + mh * 256;
                        tc::umma_f16(d, ah, bh, idesc, (kb | ks) != 0);
                        tc::umma_f16(d, al, bh, idesc, 1);
                        tc::umma_f16(d, ah, bl, idesc, 1);
                    }
                }
                tc::umma_commit(&bar->empty[stage]);
            }
            tc::umma_commit(&bar->acc_full);
        }
        __syncwarp();
    } else {
        const int quad = warp & 3, half = warp >> 2;
        tc::mbar_wait(&bar->acc_full, 0);
        tc::tcgen05_fence_after();
        float *dst0 = partial + ((long long)b * nchunks + chunk) * 256 * 256;
#pragma unroll 1
        for (int mh = 0; mh < 2; ++mh) {
            const int d = mh * 128 + quad * 32 + lane;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                const int col = half * 128 + c * 32;
                float v[32];
                tc::tmem_ld_32x32(tmem_base + ((uint32_t)(quad * 32) << 16) + mh * 256 + col, v);
                tc::tmem_ld_wait();
                float4 *dst = reinterpret_cast<float4 *>(dst0 + (long long)d * 256 + col);
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    dst[q] = make_float4(v[4 * q] * (1.f / 4096.f), v[4 * q + 1] * (1.f / 4096.f), v[4 * q + 2] * (1.f / 4096.f), v[4 * q + 3] * (1.f / 4096.f));
            }
        }
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    if (warp == 9) tc::tmem_dealloc<512>(tmem_base);
}

}  // namespace atb
}  // namespace opb

namespace opb {
namespace atb {

// per (head h, frame b): G_h = sum of chunk partials; Ksum_h = sum of column-sum partials;
// KV_h[d,:] = (G_h[d,:] Wv_h^T + Ksum_h[d] bv_h^T) / M;  P_b[o, h*64+d] = sum_q KV_h[d,q] Wm[o, h*64+q], o < n_out, written as
// split-precision planes (x64).  All folds in a fixed order.  One CTA owns a whole head, so the head's slices of WvT / WmT
// (64 KB + 64..128 KB) cross L2 -> SM once per frame-head.  Both small GEMMs are register-blocked (4 x 4 and 4 x 8 outputs per
// thread, operands as 128-bit loads: one shared-memory load per 8-11 FMAs, where the first version of this kernel needed one per 4
// and was shared-memory bound at 100-120 us per call); 512 threads at <= 64 registers, two CTAs per SM (G^T 68 KB, KV'^T 17 KB).
constexpr int kPostThreads = 512;
constexpr int kGTStride = 68;                                 // floats per row of G^T / KV'^T: 16-byte aligned rows, conflict-free 128-bit column writes
constexpr int kWRing = 2 * 2 * 16 * 64;                        // floats: 2 stages x 2 halves of the c range x 16 c x 64 q of Wv^T
constexpr int kPostSmem = (256 * kGTStride + 64 * kGTStride + 64 + kWRing) * 4 + 8 * 64 * 8;
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tc::smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__global__ void __launch_bounds__(kPostThreads, 2)
k_kv_post(const float *__restrict__ partial, int nchunks, const float2 *__restrict__ ksum_part, int nparts,
          const float *__restrict__ WvT /* [4][256 c][64 q] = Wv[h*64+q][c] */, const float *__restrict__ bv,
          const float *__restrict__ WmT /* [4][64 q][n_out o] = Wm[o][h*64+q] */, int n_out, float inv_m, float *__restrict__ ksum /* [B][256] */,
          float *__restrict__ kv_out /* [B][4][64][64] or null */, __half *__restrict__ Ph, __half *__restrict__ Pl) {
    extern __shared__ __align__(16) float post_smem[];
    float *GT = post_smem;                                                                // [256 c][68]: G^T, d contiguous
    float *KVT = post_smem + 256 * kGTStride;                                             // [64 q][68]: KV'^T, d contiguous
    float *Ks = KVT + 64 * kGTStride;                                                     // [64 d]
    double (*Kslice)[64] = reinterpret_cast<double (*)[64]>(Ks + 64);                     // [8 slices][64 d]
    float *Wring = Ks + 64 + 8 * 64 * 2;                                                  // [2 stages][2 halves][16 c][64 q]
    const int h = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int ch0 = h * 64;                                     // first K' channel of this head
    {   // G fold: thread = (column c, half of the head's 64 rows), 32 independent loads in flight per chunk; chunks summed in order
        const int c = tid & 255, d0 = (tid >> 8) * 32;
        float acc[32];
#pragma unroll
        for (int d = 0; d < 32; ++d) acc[d] = 0.f;
        for (int k = 0; k < nchunks; ++k) {
            const float *src = partial + (((long long)b * nchunks + k) * 256 + ch0 + d0) * 256 + c;
#pragma unroll
            for (int d = 0; d < 32; ++d) acc[d] += __ldg(src + d * 256);
        }
#pragma unroll
        for (int d = 0; d < 32; d += 4) *reinterpret_cast<float4 *>(&GT[c * kGTStride + d0 + d]) = make_float4(acc[d], acc[d + 1], acc[d + 2], acc[d + 3]);
    }
    {   // Ksum fold: 8 slices of the partial list per channel in float64, then the slices in order
        const int d = tid & 63, slice = tid >> 6;
        const int per = (nparts + 7) / 8, k0 = slice * per, k1 = min(nparts, k0 + per);
        double s = 0;
#pragma unroll 8
        for (int k = k0; k < k1; ++k) s += (double)__ldg(&ksum_part[((long long)b * nparts + k) * 256 + ch0 + d].x);
        Kslice[slice][d] = s;
    }
    __syncthreads();
    if (tid < 64) {
        double s = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += Kslice[k][tid];
        Ks[tid] = (float)s;
        ksum[(long long)b * 256 + ch0 + tid] = (float)s;
    }
    {   // KV'[d][q] = sum_c G[d][c] Wv[q][c]: thread = (half of the c range, 4 d, 4 q); the two halves are added in order below
        const int kh = tid >> 8, t = tid & 255, dg = t >> 4, qg = t & 15;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        // Wv^T streams through a two-stage cp.async ring, 16 c per half and stage (every thread copies 16 bytes per stage): the weights
        // come from L2, and loading them straight into registers left their latency exposed (ncu: 45 % of the samples on long_sb)
        const float *wsrc = WvT + ((long long)h * 256 + kh * 128) * 64 + 4 * t;         // + 1024 floats per stage
        float *wdst = Wring + kh * 1024 + 4 * t;                                         // + 2048 floats per stage parity
        const float *g = GT + (kh * 128) * kGTStride + 4 * dg;
        cp_async16(wdst, wsrc); cp_async_commit();
        for (int st = 0; st < 8; ++st) {
            if (st + 1 < 8) cp_async16(wdst + ((st + 1) & 1) * 2048, wsrc + (st + 1) * 1024);
            cp_async_commit();
            cp_async_wait<1>();
            __syncthreads();                                    // stage st has landed for every thread
            const float *wr = Wring + (st & 1) * 2048 + kh * 1024 + 4 * qg;
#pragma unroll
            for (int cc = 0; cc < 16; ++cc) {
                const float4 w4 = *reinterpret_cast<const float4 *>(wr + cc * 64);
                const float4 g4 = *reinterpret_cast<const float4 *>(g + (st * 16 + cc) * kGTStride);
                const float gv[4] = {g4.x, g4.y, g4.z, g4.w}, wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(gv[i], wv[j], acc[i][j]);
            }
            __syncthreads();                                    // stage st may be overwritten by the copy issued in the next iteration
        }
        __syncthreads();                                        // every read of G^T is done (and Ks is visible): its first 16 KB become scratch
        float *scr = GT + (dg * 16 + qg) * 16;                  // 16 floats per (dg, qg)
        if (kh == 1) {
#pragma unroll
            for (int i = 0; i < 4; ++i) *reinterpret_cast<float4 *>(scr + 4 * i) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        }
        __syncthreads();
        if (kh == 0) {
            const float4 bq4 = *reinterpret_cast<const float4 *>(bv + h * 64 + 4 * qg);
            const float bq[4] = {bq4.x, bq4.y, bq4.z, bq4.w};
            float v[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 o4 = *reinterpret_cast<const float4 *>(scr + 4 * i);
                const float ov[4] = {o4.x, o4.y, o4.z, o4.w};
                const float ks = Ks[4 * dg + i];
#pragma unroll
                for (int j = 0; j < 4; ++j) v[i][j] = fmaf(ks, bq[j], acc[i][j] + ov[j]) * inv_m;        // value / v_length, :93-94
                if (kv_out) *reinterpret_cast<float4 *>(kv_out + (((long long)b * 4 + h) * 64 + 4 * dg + i) * 64 + 4 * qg) = make_float4(v[i][0], v[i][1], v[i][2], v[i][3]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<float4 *>(&KVT[(4 * qg + j) * kGTStride + 4 * dg]) = make_float4(v[0][j], v[1][j], v[2][j], v[3][j]);
        }
    }
    __syncthreads();
    {   // P[o][ch0 + d] = sum_q Wm[o][h*64+q] KV'[d][q]: thread = (4 o, 8 d), 256 o per pass
        const int og = tid >> 3, dg = tid & 7;
        for (int ob = 0; ob < n_out; ob += 256) {
            const int o0 = ob + 4 * og;
            float acc[4][8];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
            // this pass's slice of Wm^T (64 q x 256 o, 64 KB) is copied into the G^T area, which is dead by now
            __syncthreads();                                    // the previous pass (or the scratch of the KV' stage) is done with it
            {
                const float *src = WmT + (long long)h * 64 * n_out + ob;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int e = (i * kPostThreads + tid) * 4, q = e >> 8, o = e & 255;
                    cp_async16(GT + e, src + (long long)q * n_out + o);
                }
                cp_async_commit();
                cp_async_wait<0>();
            }
            __syncthreads();
            const float *wm = GT + 4 * og;                                             // + 256 q
            const float *kv = KVT + 8 * dg;
#pragma unroll 4
            for (int q = 0; q < 64; ++q) {
                const float4 w4 = *reinterpret_cast<const float4 *>(wm + q * 256);
                const float4 ka = *reinterpret_cast<const float4 *>(kv + q * kGTStride), kb = *reinterpret_cast<const float4 *>(kv + q * kGTStride + 4);
                const float wv[4] = {w4.x, w4.y, w4.z, w4.w}, kk[8] = {ka.x, ka.y, ka.z, ka.w, kb.x, kb.y, kb.z, kb.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(wv[i], kk[j], acc[i][j]);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                uint32_t hh[4], ll[4];
#pragma unroll
                for (int j = 0; j < 8; j += 2) {
                    __half h0, l0, h1, l1;
                    tc::split_f16(acc[i][j] * 64.f, h0, l0); tc::split_f16(acc[i][j + 1] * 64.f, h1, l1);
                    hh[j / 2] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
                    ll[j / 2] = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
                }
                const long long e = ((long long)b * n_out + o0 + i) * 256 + ch0 + 8 * dg;
                *reinterpret_cast<uint4 *>(Ph + e) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
                *reinterpret_cast<uint4 *>(Pl + e) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
            }
        }
    }
}

}  // namespace atb
}  // namespace opb

extern "C" int op_atb_chunks(int M) { return opb::ceil_div(M > 0 ? M : 1, opb::atb::kChunk); }

extern "C" int op_atb_tc(const void *Ah, const void *Al, const void *Bh, const void *Bl, int B, int M, float *partial, void *stream) {
    using namespace opb;
    OPB_CHECK_ARG(Ah && Al && Bh && Bl && partial, "op_atb_tc: null pointer");
    OPB_CHECK_ARG(B > 0 && M > 0 && B <= 65535, "op_atb_tc: bad shape B=%d M=%d", B, M);
    CUtensorMap tah, tal, tbh, tbl;
    int rc;
    if ((rc = make_tmap_f16_3d(&tah, Ah, B, M, kD, atb::KB)) || (rc = make_tmap_f16_3d(&tal, Al, B, M, kD, atb::KB)) ||
        (rc = make_tmap_f16_3d(&tbh, Bh, B, M, kD, atb::KB)) || (rc = make_tmap_f16_3d(&tbl, Bl, B, M, kD, atb::KB)))
        return rc;
    static bool configured = false;
    if (!configured) {
        OPB_CHECK_CUDA(cudaFuncSetAttribute(atb::k_atb_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, atb::SMEM_BYTES));
        configured = true;
    }
    const int nchunks = ceil_div(M, atb::kChunk);
    atb::k_atb_tc<<<dim3(nchunks, B), atb::THREADS, atb::SMEM_BYTES, (cudaStream_t)stream>>>(tah, tal, tbh, tbl, M, nchunks, partial);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_kv_post(const float *partial, int nchunks, const float *ksum_partial, int nparts, const float *WvT, const float *bv,
                          const float *WmT, int n_out, int B, float inv_m, float *ksum, float *kv_out, void *Ph, void *Pl, void *stream) {
    using namespace opb;
    OPB_CHECK_ARG(partial && ksum_partial && WvT && bv && WmT && ksum && Ph && Pl, "op_kv_post: null pointer");
    OPB_CHECK_ARG(B > 0 && nchunks > 0 && nparts > 0 && B <= 65535 && n_out > 0 && n_out % 256 == 0, "op_kv_post: bad shape (n_out=%d)", n_out);
    static bool post_configured = false;
    if (!post_configured) {
        OPB_CHECK_CUDA(cudaFuncSetAttribute(atb::k_kv_post, cudaFuncAttributeMaxDynamicSharedMemorySize, atb::kPostSmem));
        post_configured = true;
    }
    atb::k_kv_post<<<dim3(4, B), atb::kPostThreads, atb::kPostSmem, (cudaStream_t)stream>>>(partial, nchunks, (const float2 *)ksum_partial, nparts, WvT, bv, WmT,
                                                                            n_out, inv_m, ksum, kv_out, (__half *)Ph, (__half *)Pl);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
