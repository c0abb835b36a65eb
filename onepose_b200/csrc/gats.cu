// (1) GATs leaf-aggregation graph attention -- reference GATs.py:62-123.
//
// Bandwidth kernel.  One warp owns one 3D point n: its L leaf descriptors (L x 256 fp32) are staged
// once in shared memory with 128-bit loads and reused for every frame b of the batch (the leaves are
// a static per-object DB, only h3 changes per frame), so the algorithmic traffic per frame is
// 2 x 1 KB per point plus 1 KB x L / B.  Dot products and the (1+L)-way softmax are warp-shuffle
// reductions; lane j keeps the logit of leaf j.
#include "common.cuh"

namespace opb {

constexpr int kGatsWarps = 4;

__device__ __forceinline__ float dot8(const float4 &a0, const float4 &a1, const float4 &b0, const float4 &b1) {
    return a0.x * b0.x + a0.y * b0.y + a0.z * b0.z + a0.w * b0.w + a1.x * b1.x + a1.y * b1.y + a1.z * b1.z + a1.w * b1.w;
}

__global__ void __launch_bounds__(kGatsWarps * 32)
k_gats_leaf_aggregate(const float *__restrict__ leaf, long long leaf_bstride, const float *h3L,
                      const float *h3_add, const float *__restrict__ v1, const float *__restrict__ v2,
                      float *out, int B, int N3, int L, float alpha, int include_self, int apply_elu) {
    extern __shared__ float4 smem4[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.x * kGatsWarps + warp;
    if (n >= N3) return;                       // warp-uniform; only __syncwarp below
    float4 *sl = smem4 + (size_t)warp * L * 64;   // L x 256 floats = L x 64 float4
    const float4 *v1v = reinterpret_cast<const float4 *>(v1), *v2v = reinterpret_cast<const float4 *>(v2);
    const float4 a0 = v1v[lane], a1 = v1v[lane + 32], c0 = v2v[lane], c1 = v2v[lane + 32];
    float s_leaf = 0.f;                        // lane j: logit part of leaf j
    for (int b = 0; b < B; ++b) {
        if (b == 0 || leaf_bstride != 0) {
            const float4 *lp = reinterpret_cast<const float4 *>(leaf + (long long)b * leaf_bstride + (long long)n * L * kD);
            __syncwarp();
            for (int j = 0; j < L; ++j) {
                float4 x0 = __ldg(lp + j * 64 + lane), x1 = __ldg(lp + j * 64 + 32 + lane);
                sl[j * 64 + lane] = x0;
                sl[j * 64 + 32 + lane] = x1;
                float d = warp_sum(dot8(x0, x1, a0, a1));
                if (lane == j) s_leaf = d;
            }
            __syncwarp();
        }
        const long long row = ((long long)b * N3 + n) * (kD / 4);
        const float4 *hp = reinterpret_cast<const float4 *>(h3L) + row;
        const float4 h0 = hp[lane], h1 = hp[lane + 32];
        const float t = warp_sum(dot8(h0, h1, c0, c1));
        // the self entry re-uses the a[256:] projection (GATs.py:115-120: cat([wh_3d_, wh_2d_])), so e_self = t + t
        float e_self = t + t;
        e_self = e_self > 0.f ? e_self : alpha * e_self;
        float e = t + s_leaf;
        e = e > 0.f ? e : alpha * e;
        if (lane >= L) e = -INFINITY;
        float m = warp_max(e);
        if (include_self) m = fmaxf(m, e_self);
        float p = lane < L ? expf(e - m) : 0.f;
        const float p_self = include_self ? expf(e_self - m) : 0.f;
        const float denom = warp_sum(p) + p_self;
        p = p / denom;
        const float w_self = p_self / denom;
        float4 o0 = make_float4(w_self * h0.x, w_self * h0.y, w_self * h0.z, w_self * h0.w);
        float4 o1 = make_float4(w_self * h1.x, w_self * h1.y, w_self * h1.z, w_self * h1.w);
        for (int j = 0; j < L; ++j) {
            const float a = __shfl_sync(0xffffffffu, p, j);
            const float4 x0 = sl[j * 64 + lane], x1 = sl[j * 64 + 32 + lane];
            o0.x += a * x0.x; o0.y += a * x0.y; o0.z += a * x0.z; o0.w += a * x0.w;
            o1.x += a * x1.x; o1.y += a * x1.y; o1.z += a * x1.z; o1.w += a * x1.w;
        }
        if (!include_self) {      // GATs.py:97-99
            o0.x = o0.x * 0.5f + h0.x; o0.y = o0.y * 0.5f + h0.y; o0.z = o0.z * 0.5f + h0.z; o0.w = o0.w * 0.5f + h0.w;
            o1.x = o1.x * 0.5f + h1.x; o1.y = o1.y * 0.5f + h1.y; o1.z = o1.z * 0.5f + h1.z; o1.w = o1.w * 0.5f + h1.w;
        }
        if (h3_add) {             // GATs.py:93-94
            const float4 *ap = reinterpret_cast<const float4 *>(h3_add) + row;
            const float4 r0 = ap[lane], r1 = ap[lane + 32];      // may alias out: plain loads
            o0.x += r0.x; o0.y += r0.y; o0.z += r0.z; o0.w += r0.w;
            o1.x += r1.x; o1.y += r1.y; o1.z += r1.z; o1.w += r1.w;
        }
        if (apply_elu) {
            o0.x = elu(o0.x); o0.y = elu(o0.y); o0.z = elu(o0.z); o0.w = elu(o0.w);
            o1.x = elu(o1.x); o1.y = elu(o1.y); o1.z = elu(o1.z); o1.w = elu(o1.w);
        }
        float4 *op = reinterpret_cast<float4 *>(out) + row;
        op[lane] = o0;
        op[lane + 32] = o1;
    }
}

}  // namespace opb

extern "C" int op_gats_leaf_aggregate(const float *leaf, long long leaf_bstride, const float *h3L, const float *h3_add,
                                      const float *v1, const float *v2, float *out, int B, int N3, int L,
                                      float alpha, int include_self, int apply_elu, void *stream) {
    OPB_CHECK_ARG(leaf && h3L && v1 && v2 && out, "op_gats_leaf_aggregate: null pointer");
    OPB_CHECK_ARG(B >= 0 && N3 >= 0, "op_gats_leaf_aggregate: bad shape B=%d N3=%d", B, N3);
    OPB_CHECK_ARG(L >= 1 && L <= 32, "op_gats_leaf_aggregate: num_leaf=%d not in [1,32]", L);
    if (B == 0 || N3 == 0) return OPB_OK;
    const size_t smem = (size_t)opb::kGatsWarps * L * opb::kD * sizeof(float);
    static thread_local size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        OPB_CHECK_CUDA(cudaFuncSetAttribute(opb::k_gats_leaf_aggregate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    const int grid = opb::ceil_div(N3, opb::kGatsWarps);
    opb::k_gats_leaf_aggregate<<<grid, opb::kGatsWarps * 32, smem, (cudaStream_t)stream>>>(
        leaf, leaf_bstride, h3L, h3_add, v1, v2, out, B, N3, L, alpha, include_self, apply_elu);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
