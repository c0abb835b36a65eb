// (2) linear attention pieces -- reference linear_attention, GATs_SuperGlue.py:88-99.
//
// Channels are head-major here (c = h*64 + d); the host permutes the projection weights once so the
// reference's interleaved `.view(b, 64, 4, n)` split (channel c -> d = c//4, h = c%4, :116) becomes
// contiguous 64-wide heads.  The KV reduction over the source points is a grid-wide reduction: each
// block reduces a chunk of points and a second kernel folds the chunk partials in a fixed order, so
// results are bit-reproducible run to run.
#include "common.cuh"
#include "tc_common.cuh"

namespace opb {

constexpr int kKvChunk = 256;     // source points per block
constexpr int kKvTile = 32;       // points staged per smem tile
constexpr int kKvPartial = kDh * kDh + kDh;   // 64x64 KV + 64 Ksum

__global__ void __launch_bounds__(256)
k_kv_partial(const float *__restrict__ kv, float *__restrict__ partial, int M, int nchunks) {
    __shared__ __align__(16) float Ks[kKvTile][kDh];
    __shared__ __align__(16) float Vs[kKvTile][kDh];
    const int chunk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int tid = threadIdx.x, qx = tid & 15, dy = tid >> 4;
    const float fM = (float)M;
    float acc[4][4] = {};
    float ksum[4] = {};
    const int m_begin = chunk * kKvChunk, m_end = min(M, m_begin + kKvChunk);
    for (int m0 = m_begin; m0 < m_end; m0 += kKvTile) {
        // 32 points x 16 float4 for K and for V: 512 float4 each, 2 per thread each
#pragma unroll
        for (int it = 0; it < 2; ++it) {
            const int idx = tid + it * 256, r = idx >> 4, c4 = idx & 15;
            const int m = m0 + r;
            float4 kq = make_float4(0, 0, 0, 0), vq = kq;
            if (m < m_end) {
                const float *row = kv + ((long long)b * M + m) * (2 * kD);
                kq = *reinterpret_cast<const float4 *>(row + h * kDh + c4 * 4);
                vq = *reinterpret_cast<const float4 *>(row + kD + h * kDh + c4 * 4);
                vq.x /= fM; vq.y /= fM; vq.z /= fM; vq.w /= fM;       // value / v_length (:93-94)
            }
            *reinterpret_cast<float4 *>(&Ks[r][c4 * 4]) = kq;
            *reinterpret_cast<float4 *>(&Vs[r][c4 * 4]) = vq;
        }
        __syncthreads();
#pragma unroll 8
        for (int r = 0; r < kKvTile; ++r) {
            const float4 a = *reinterpret_cast<const float4 *>(&Ks[r][dy * 4]);
            const float4 v = *reinterpret_cast<const float4 *>(&Vs[r][qx * 4]);
            const float aa[4] = {a.x, a.y, a.z, a.w}, vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                ksum[i] += aa[i];
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(aa[i], vv[j], acc[i][j]);
            }
        }
        __syncthreads();
    }
    float *dst = partial + (((long long)b * kHeads + h) * nchunks + chunk) * kKvPartial;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        *reinterpret_cast<float4 *>(dst + (dy * 4 + i) * kDh + qx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        if (qx == 0) dst[kDh * kDh + dy * 4 + i] = ksum[i];
    }
}

__global__ void k_kv_finalize(const float *__restrict__ partial, float *__restrict__ KV, float *__restrict__ Ksum, int nchunks) {
    const int bh = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= kKvPartial) return;
    float s = 0.f;
    for (int c = 0; c < nchunks; ++c) s += partial[((long long)bh * nchunks + c) * kKvPartial + e];
    if (e < kDh * kDh) KV[(long long)bh * kDh * kDh + e] = s;
    else Ksum[(long long)bh * kDh + (e - kDh * kDh)] = s;
}

// out[n, h*64+q] = (Q'[n,h,:] . KV[h,:,q]) * M / (Q'[n,h,:] . Ksum[h,:] + 1e-6)
__global__ void __launch_bounds__(256)
k_attn_apply(const float *__restrict__ Qp, const float *__restrict__ KV, const float *__restrict__ Ksum,
             float *__restrict__ out, __half *__restrict__ out_hi, __half *__restrict__ out_lo, int N, int M) {
    __shared__ __align__(16) float KVs[kDh][kDh];
    __shared__ __align__(16) float Qs[64][kDh + 4];
    __shared__ float Kss[kDh];
    const int n0 = blockIdx.x * 64, h = blockIdx.y, b = blockIdx.z;
    const int tid = threadIdx.x, qx = tid & 15, py = tid >> 4;
    const float *kvsrc = KV + ((long long)b * kHeads + h) * kDh * kDh;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        const int idx = tid + it * 256, r = idx >> 4, c4 = idx & 15;
        *reinterpret_cast<float4 *>(&KVs[r][c4 * 4]) = *reinterpret_cast<const float4 *>(kvsrc + r * kDh + c4 * 4);
        const int n = n0 + r;
        float4 q = make_float4(0, 0, 0, 0);
        if (n < N) q = *reinterpret_cast<const float4 *>(Qp + ((long long)b * N + n) * kD + h * kDh + c4 * 4);
        *reinterpret_cast<float4 *>(&Qs[r][c4 * 4]) = q;
    }
    if (tid < kDh) Kss[tid] = Ksum[((long long)b * kHeads + h) * kDh + tid];
    __syncthreads();
    float acc[4][4] = {};
    float den[4] = {};
#pragma unroll 8
    for (int d = 0; d < kDh; ++d) {
        const float4 v = *reinterpret_cast<const float4 *>(&KVs[d][qx * 4]);
        const float ks = Kss[d];
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float q = Qs[py * 4 + i][d];
            den[i] = fmaf(q, ks, den[i]);
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(q, vv[j], acc[i][j]);
        }
    }
    const float fM = (float)M;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = n0 + py * 4 + i;
        if (n >= N) continue;
        const float z = 1.f / (den[i] + 1e-6f);
        float4 o = make_float4(acc[i][0] * z * fM, acc[i][1] * z * fM, acc[i][2] * z * fM, acc[i][3] * z * fM);
        const long long off = ((long long)b * N + n) * kD + h * kDh + qx * 4;
        if (out) *reinterpret_cast<float4 *>(out + off) = o;
        if (out_hi) {            // split-precision planes of 64*x for the tcgen05 merge projection (linear_tc.cuh)
            __half hh[4], ll[4];
            tc::split_f16(o.x * 64.f, hh[0], ll[0]); tc::split_f16(o.y * 64.f, hh[1], ll[1]);
            tc::split_f16(o.z * 64.f, hh[2], ll[2]); tc::split_f16(o.w * 64.f, hh[3], ll[3]);
            *reinterpret_cast<uint2 *>(out_hi + off) = *reinterpret_cast<uint2 *>(hh);
            *reinterpret_cast<uint2 *>(out_lo + off) = *reinterpret_cast<uint2 *>(ll);
        }
    }
}

}  // namespace opb

extern "C" size_t op_kv_reduce_workspace_bytes(int B, int M) {
    const size_t nchunks = (size_t)opb::ceil_div(M > 0 ? M : 1, opb::kKvChunk);
    return (size_t)(B > 0 ? B : 1) * opb::kHeads * nchunks * opb::kKvPartial * sizeof(float);
}

extern "C" int op_kv_reduce(const float *kv, float *KV, float *Ksum, int B, int M, void *workspace,
                            size_t workspace_bytes, void *stream) {
    OPB_CHECK_ARG(kv && KV && Ksum && workspace, "op_kv_reduce: null pointer");
    OPB_CHECK_ARG(B > 0 && M > 0 && B <= 65535, "op_kv_reduce: bad shape B=%d M=%d", B, M);
    if (workspace_bytes < op_kv_reduce_workspace_bytes(B, M)) {
        opb::set_error("op_kv_reduce: workspace %zu < %zu bytes", workspace_bytes, op_kv_reduce_workspace_bytes(B, M));
        return OPB_ERR_WORKSPACE;
    }
    const int nchunks = opb::ceil_div(M, opb::kKvChunk);
    dim3 grid(nchunks, opb::kHeads, B);
    opb::k_kv_partial<<<grid, 256, 0, (cudaStream_t)stream>>>(kv, (float *)workspace, M, nchunks);
    OPB_CHECK_LAUNCH();
    dim3 grid2(opb::ceil_div(opb::kKvPartial, 256), B * opb::kHeads);
    opb::k_kv_finalize<<<grid2, 256, 0, (cudaStream_t)stream>>>((const float *)workspace, KV, Ksum, nchunks);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_attn_apply(const float *Qp, const float *KV, const float *Ksum, float *out, int B, int N, int M,
                             void *stream) {
    OPB_CHECK_ARG(Qp && KV && Ksum && out, "op_attn_apply: null pointer");
    OPB_CHECK_ARG(B > 0 && N > 0 && M > 0 && B <= 65535, "op_attn_apply: bad shape B=%d N=%d M=%d", B, N, M);
    dim3 grid(opb::ceil_div(N, 64), opb::kHeads, B);
    opb::k_attn_apply<<<grid, 256, 0, (cudaStream_t)stream>>>(Qp, KV, Ksum, out, nullptr, nullptr, N, M);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_attn_apply_planes(const float *Qp, const float *KV, const float *Ksum, void *out_hi, void *out_lo, int B, int N,
                                    int M, void *stream) {
    OPB_CHECK_ARG(Qp && KV && Ksum && out_hi && out_lo, "op_attn_apply_planes: null pointer");
    OPB_CHECK_ARG(B > 0 && N > 0 && M > 0 && B <= 65535, "op_attn_apply_planes: bad shape B=%d N=%d M=%d", B, N, M);
    dim3 grid(opb::ceil_div(N, 64), opb::kHeads, B);
    opb::k_attn_apply<<<grid, 256, 0, (cudaStream_t)stream>>>(Qp, KV, Ksum, nullptr, (__half *)out_hi, (__half *)out_lo, N, M);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

// ---- op_kv_reduce_static: the same KV reduction for callers that may not allocate -----------------------------------------
// No counterpart in the reference; added (callers under CUDA graph capture, tight serving
// loops and harnesses that forbid cudaMalloc cannot provide a workspace).  The chunk partials live in a file-scope device array of
// OPB_KV_STATIC_SCRATCH_BYTES (8 MiB) that every GPU loading the library holds until the process exits.
//  * In HOST code the array's name is only a stand-in for the symbol, not a device address: the address is resolved once per
//    device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels get.
//  * Nothing in the array has to start at zero: k_kv_partial writes every element k_kv_finalize reads (all nchunks partials of
//    every (frame, head)), so the previous call's contents are never observed and nothing is cleared.
//  * All calls on a device share the one array: two calls that overlap on the GPU (different streams, or graph branches without
//    a dependency) race on it.  Calls in one stream are ordered and safe.
// The entry point calls no cudaMalloc / cudaFree and no allocating library routine.
#include <mutex>

namespace opb {

constexpr size_t kKvStaticFloats = OPB_KV_STATIC_SCRATCH_BYTES / sizeof(float);
static __device__ __align__(256) float g_kv_static_scratch[kKvStaticFloats];

// device address of g_kv_static_scratch on the CURRENT device (resolved at first use per device, then cached)
static cudaError_t kv_static_scratch_ptr(float **out) {
    constexpr int kMaxDevices = 64;
    static float *cached[kMaxDevices] = {};
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        float *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_kv_static_scratch);     // `dev` is current: this is that device's copy of the array
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        cached[dev] = p;
    }
    *out = cached[dev];
    return cudaSuccess;
}

}  // namespace opb

extern "C" size_t op_kv_reduce_static_max_bytes(void) { return OPB_KV_STATIC_SCRATCH_BYTES; }

extern "C" int op_kv_reduce_static(const float *kv, float *KV, float *Ksum, int B, int M, void *stream) {
    OPB_CHECK_ARG(kv && KV && Ksum, "op_kv_reduce_static: null pointer");
    OPB_CHECK_ARG(B > 0 && M > 0 && B <= 65535, "op_kv_reduce_static: bad shape B=%d M=%d", B, M);
    const size_t need = op_kv_reduce_workspace_bytes(B, M);
    if (need > (size_t)OPB_KV_STATIC_SCRATCH_BYTES) {
        opb::set_error("op_kv_reduce_static: B=%d M=%d needs %zu bytes of scratch, the static array holds %zu (B * ceil(M / 256) <= %zu); "
                       "use op_kv_reduce with a caller-owned workspace", B, M, need, (size_t)OPB_KV_STATIC_SCRATCH_BYTES,
                       (size_t)OPB_KV_STATIC_SCRATCH_BYTES / (opb::kHeads * opb::kKvPartial * sizeof(float)));
        return OPB_ERR_WORKSPACE;
    }
    float *scratch = nullptr;
    OPB_CHECK_CUDA(opb::kv_static_scratch_ptr(&scratch));
    const int nchunks = opb::ceil_div(M, opb::kKvChunk);
    dim3 grid(nchunks, opb::kHeads, B);
    opb::k_kv_partial<<<grid, 256, 0, (cudaStream_t)stream>>>(kv, scratch, M, nchunks);
    OPB_CHECK_LAUNCH();
    dim3 grid2(opb::ceil_div(opb::kKvPartial, 256), B * opb::kHeads);
    opb::k_kv_finalize<<<grid2, 256, 0, (cudaStream_t)stream>>>(scratch, KV, Ksum, nchunks);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
