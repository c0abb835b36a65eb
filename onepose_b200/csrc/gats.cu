// (1) GATs leaf-aggregation graph attention -- reference GATs.py:62-123.
//
// Bandwidth kernel.  One warp owns one 3D point n: its L leaf descriptors (L x 256 fp32) are staged
// once in shared memory with 128-bit loads and reused for every frame b of the batch (the leaves are
// a static per-object DB, only h3 changes per frame), so the algorithmic traffic per frame is
// 2 x 1 KB per point plus 1 KB x L / B.  Dot products and the (1+L)-way softmax are warp-shuffle
// reductions; lane j keeps the logit of leaf j.
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"

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

// The same layer on the split-precision planes of the tensor-core path (linear_tc.cuh), in place, for h3L == the 3D stream
// itself (with_linear_transform = False): no fp32 copy of the stream exists any more.  kGatsPW points per block, kGatsFW warps
// per point sharing its staged leaves and taking alternate frames; the next two rows of a warp are in flight while one is
// being processed; the column sums of the output (the analytic InstanceNorm mean of the next MLP needs them) are reduced over
// the block's points per frame and written as partials [B][blocks][256] (sum, 0) -- one __syncthreads per kGatsFW frames,
// double-buffered.
constexpr int kGatsPW = 8, kGatsFW = 2;
static_assert(kGatsPW * 32 == kD, "the column-sum fold maps one thread to one channel");

__device__ __forceinline__ void planes_to_f4(const uint2 &h, const uint2 &l, float4 &o) {
    const float2 h0 = __half22float2(*reinterpret_cast<const __half2 *>(&h.x)), h1 = __half22float2(*reinterpret_cast<const __half2 *>(&h.y));
    const float2 l0 = __half22float2(*reinterpret_cast<const __half2 *>(&l.x)), l1 = __half22float2(*reinterpret_cast<const __half2 *>(&l.y));
    o = make_float4((h0.x + l0.x) * (1.f / 64.f), (h0.y + l0.y) * (1.f / 64.f), (h1.x + l1.x) * (1.f / 64.f), (h1.y + l1.y) * (1.f / 64.f));
}
// packed conversions (cvt.rn.f16x2.f32): the same values as tc::split_f16 element by element, three fifths of the instructions
__device__ __forceinline__ void f4_to_planes(const float4 &v, uint2 &h, uint2 &l) {
    const float s0 = v.x * 64.f, s1 = v.y * 64.f, s2 = v.z * 64.f, s3 = v.w * 64.f;
    const __half2 h01 = __floats2half2_rn(s0, s1), h23 = __floats2half2_rn(s2, s3);
    const float2 f01 = __half22float2(h01), f23 = __half22float2(h23);
    const __half2 l01 = __floats2half2_rn(s0 - f01.x, s1 - f01.y), l23 = __floats2half2_rn(s2 - f23.x, s3 - f23.y);
    h = make_uint2(*reinterpret_cast<const uint32_t *>(&h01), *reinterpret_cast<const uint32_t *>(&h23));
    l = make_uint2(*reinterpret_cast<const uint32_t *>(&l01), *reinterpret_cast<const uint32_t *>(&l23));
}
// ELU through MUFU.EX2 (absolute error ~2e-7, below the 22-bit planes the result is stored in)
__device__ __forceinline__ float elu_fast(float x) { return x > 0.f ? x : tc::ex2(x * 1.4426950408889634f) - 1.f; }

// LT > 0: the number of leaves is a compile-time constant (8 and 16, the shipped object models), so the loops over the leaves are
// unrolled four at a time without a remainder loop (the 64-register budget of two 512-thread blocks per SM does not allow more);
// LT = 0 takes L at run time.
// FL >= 0: include_self (bit 0) and add_self (bit 1) are compile-time constants too (FL = 1 is the shipped layer).
// REG = 1 (LT = 8, one DB shared by all frames): the point's leaves live in REGISTERS (8 leaves x 8 channels per lane = 64) for the whole
// frame loop instead of being re-read from shared memory for every frame -- ncu of the shared-memory form at B = 64: l1tex 83 % of peak,
// top stall short_scoreboard, DRAM at 0.56 of the copy peak.  128 registers: one 512-thread block per SM.
template <int LT, int FL, int REG = 0>
__global__ void __launch_bounds__(kGatsPW * kGatsFW * 32, REG ? 1 : 2)
k_gats_planes(const float *__restrict__ leaf, long long leaf_bstride, __half *xh, __half *xl, const float *__restrict__ v1,
              const float *__restrict__ v2, float2 *__restrict__ colsum, int B, int N3, int Lrt, float alpha, int include_self_rt,
              int add_self_rt) {
    const int L = LT > 0 ? LT : Lrt;
    const bool include_self = FL >= 0 ? (FL & 1) != 0 : include_self_rt != 0;
    const bool add_self = FL >= 0 ? (FL & 2) != 0 : add_self_rt != 0;
    extern __shared__ float4 smem4[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pt = warp % kGatsPW, fw = warp / kGatsPW;       // point of the block, frame way
    const int n = blockIdx.x * kGatsPW + pt;
    const bool valid = n < N3;                 // warp-uniform; invalid warps only take part in the barriers
    float4 *sl = smem4 + (size_t)pt * L * 64;
    float4 *red = smem4 + (size_t)kGatsPW * L * 64;          // [2][kGatsFW][kGatsPW][64] float4
    const float4 *v1v = reinterpret_cast<const float4 *>(v1), *v2v = reinterpret_cast<const float4 *>(v2);
    const float4 c0 = v2v[lane], c1 = v2v[lane + 32];
    // lane l owns channels 4l..4l+3 and 128+4l..128+4l+3 (the float4 pattern of the staged leaves)
    auto load_row = [&](int b, uint2 (&r)[4]) {
        const long long e = ((long long)b * N3 + n) * kD + 4 * lane;
        r[0] = *reinterpret_cast<const uint2 *>(xh + e); r[1] = *reinterpret_cast<const uint2 *>(xh + e + 128);
        r[2] = *reinterpret_cast<const uint2 *>(xl + e); r[3] = *reinterpret_cast<const uint2 *>(xl + e + 128);
    };
    uint2 cur[4] = {}, nx1[4] = {}, nx2[4] = {};
    if (valid) { if (fw < B) load_row(fw, cur); if (fw + kGatsFW < B) load_row(fw + kGatsFW, nx1); }
    float s_leaf = 0.f;                        // lane j: logit part of leaf j
    auto stage_leaves = [&](int b) {           // the warps of one point each stage 1/kGatsFW of its leaves
        const float4 *lp = reinterpret_cast<const float4 *>(leaf + (long long)b * leaf_bstride + (long long)n * L * kD);
        for (int j = fw; j < L; j += kGatsFW) {
            sl[j * 64 + lane] = __ldg(lp + j * 64 + lane);
            sl[j * 64 + 32 + lane] = __ldg(lp + j * 64 + 32 + lane);
        }
    };
    auto leaf_logits = [&]() {
        const float4 a0 = v1v[lane], a1 = v1v[lane + 32];
        for (int j = 0; j < L; ++j) {
            const float d = warp_sum(dot8(sl[j * 64 + lane], sl[j * 64 + 32 + lane], a0, a1));
            if (lane == j) s_leaf = d;
        }
    };
    if (leaf_bstride == 0) {                   // one DB for every frame: staged once
        if (valid) stage_leaves(0);
        __syncthreads();
        if (valid) leaf_logits();
    }
    float4 lf[REG ? LT : 1][2];
    if (REG && valid) {
#pragma unroll
        for (int j = 0; j < (REG ? LT : 1); ++j) { lf[j][0] = sl[j * 64 + lane]; lf[j][1] = sl[j * 64 + 32 + lane]; }
    }
    const int steps = (B + kGatsFW - 1) / kGatsFW;
    for (int it = 0; it < steps; ++it) {
        const int b = it * kGatsFW + fw;
        const bool live = valid && b < B;
        if (live && b + 2 * kGatsFW < B) load_row(b + 2 * kGatsFW, nx2);
        float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0;
        auto process = [&]() {
            float4 h0, h1;
            planes_to_f4(cur[0], cur[2], h0); planes_to_f4(cur[1], cur[3], h1);
            const float t = warp_sum(dot8(h0, h1, c0, c1));
            float e_self = t + t;              // GATs.py:115-120: the self entry re-uses the a[256:] projection
            e_self = e_self > 0.f ? e_self : alpha * e_self;
            float e = t + s_leaf;
            e = e > 0.f ? e : alpha * e;
            if (lane >= L) e = -INFINITY;
            float m = warp_max(e);
            if (include_self) m = fmaxf(m, e_self);
            // softmax weights through MUFU.EX2 and one reciprocal (2 ulp; the result is stored in 22-bit planes)
            float pr = lane < L ? tc::ex2((e - m) * 1.4426950408889634f) : 0.f;
            const float p_self = include_self ? tc::ex2((e_self - m) * 1.4426950408889634f) : 0.f;
            const float inv = 1.f / (warp_sum(pr) + p_self);
            pr = pr * inv;
            const float w_self = p_self * inv;
            o0 = make_float4(w_self * h0.x, w_self * h0.y, w_self * h0.z, w_self * h0.w);
            o1 = make_float4(w_self * h1.x, w_self * h1.y, w_self * h1.z, w_self * h1.w);
            if (REG) {
#pragma unroll
                for (int j = 0; j < (REG ? LT : 1); ++j) {
                    const float a = __shfl_sync(0xffffffffu, pr, j);
                    const float4 x0 = lf[j][0], x1 = lf[j][1];
                    o0.x += a * x0.x; o0.y += a * x0.y; o0.z += a * x0.z; o0.w += a * x0.w;
                    o1.x += a * x1.x; o1.y += a * x1.y; o1.z += a * x1.z; o1.w += a * x1.w;
                }
            } else {
#pragma unroll 4
                for (int j = 0; j < L; ++j) {
                    const float a = __shfl_sync(0xffffffffu, pr, j);
                    const float4 x0 = sl[j * 64 + lane], x1 = sl[j * 64 + 32 + lane];
                    o0.x += a * x0.x; o0.y += a * x0.y; o0.z += a * x0.z; o0.w += a * x0.w;
                    o1.x += a * x1.x; o1.y += a * x1.y; o1.z += a * x1.z; o1.w += a * x1.w;
                }
            }
            if (!include_self) {               // GATs.py:97-99
                o0.x = o0.x * 0.5f + h0.x; o0.y = o0.y * 0.5f + h0.y; o0.z = o0.z * 0.5f + h0.z; o0.w = o0.w * 0.5f + h0.w;
                o1.x = o1.x * 0.5f + h1.x; o1.y = o1.y * 0.5f + h1.y; o1.z = o1.z * 0.5f + h1.z; o1.w = o1.w * 0.5f + h1.w;
            }
            if (add_self) {                    // GATs.py:93-94
                o0.x += h0.x; o0.y += h0.y; o0.z += h0.z; o0.w += h0.w;
                o1.x += h1.x; o1.y += h1.y; o1.z += h1.z; o1.w += h1.w;
            }
            o0.x = elu_fast(o0.x); o0.y = elu_fast(o0.y); o0.z = elu_fast(o0.z); o0.w = elu_fast(o0.w);
            o1.x = elu_fast(o1.x); o1.y = elu_fast(o1.y); o1.z = elu_fast(o1.z); o1.w = elu_fast(o1.w);
            uint2 oh0, ol0, oh1, ol1;
            f4_to_planes(o0, oh0, ol0); f4_to_planes(o1, oh1, ol1);
            const long long e_out = ((long long)b * N3 + n) * kD + 4 * lane;
            *reinterpret_cast<uint2 *>(xh + e_out) = oh0; *reinterpret_cast<uint2 *>(xh + e_out + 128) = oh1;
            *reinterpret_cast<uint2 *>(xl + e_out) = ol0; *reinterpret_cast<uint2 *>(xl + e_out + 128) = ol1;
        };
        if (leaf_bstride == 0) {
            if (live) process();
        } else {                               // per-frame DB copies (reference pack_data style): the point's warps take turns
            for (int f = 0; f < kGatsFW; ++f) {
                __syncthreads();               // everybody is done with the previous frame's leaves
                if (valid && it * kGatsFW + f < B) stage_leaves(it * kGatsFW + f);
                __syncthreads();
                if (live && f == fw) { leaf_logits(); process(); }
            }
        }
        float4 *rb = red + (((size_t)(it & 1) * kGatsFW + fw) * kGatsPW + pt) * 64;
        rb[lane] = o0; rb[lane + 32] = o1;
        __syncthreads();
        {
            const int f = threadIdx.x >> 8, ch = threadIdx.x & 255, bb = it * kGatsFW + f;
            if (bb < B) {
                const float *rf = reinterpret_cast<const float *>(red + ((size_t)(it & 1) * kGatsFW + f) * kGatsPW * 64);
                float s = 0.f;
#pragma unroll
                for (int w = 0; w < kGatsPW; ++w) s += rf[w * 256 + ch];
                colsum[((long long)bb * gridDim.x + blockIdx.x) * 256 + ch] = make_float2(s, 0.f);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) { cur[i] = nx1[i]; nx1[i] = nx2[i]; }
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

extern "C" int op_gats_planes_parts(int N3) { return opb::ceil_div(N3 > 0 ? N3 : 1, opb::kGatsPW); }

extern "C" int op_gats_leaf_aggregate_planes(const float *leaf, long long leaf_bstride, void *xh, void *xl, const float *v1, const float *v2,
                                             float *colsum_partial, int B, int N3, int L, float alpha, int include_self, int add_self,
                                             void *stream) {
    OPB_CHECK_ARG(leaf && xh && xl && v1 && v2 && colsum_partial, "op_gats_leaf_aggregate_planes: null pointer");
    OPB_CHECK_ARG(B >= 0 && N3 >= 0, "op_gats_leaf_aggregate_planes: bad shape B=%d N3=%d", B, N3);
    OPB_CHECK_ARG(L >= 1 && L <= 24, "op_gats_leaf_aggregate_planes: num_leaf=%d not in [1,24]", L);   // leaves + fold buffers within 227 KB
    if (B == 0 || N3 == 0) return OPB_OK;
    const size_t smem = ((size_t)opb::kGatsPW * L + 2 * opb::kGatsFW * opb::kGatsPW) * opb::kD * sizeof(float);
    auto launch = [&](auto kern) -> cudaError_t {
        if (smem > 48 * 1024) {                       // cheap; the limit is per kernel instantiation
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        const int grid = opb::ceil_div(N3, opb::kGatsPW);
        kern<<<grid, opb::kGatsPW * opb::kGatsFW * 32, smem, (cudaStream_t)stream>>>(leaf, leaf_bstride, (__half *)xh, (__half *)xl, v1, v2,
                                                                                 (float2 *)colsum_partial, B, N3, L, alpha, include_self, add_self);
        return cudaSuccess;
    };
    const bool shipped = include_self && !add_self;
    // OPB_GATS_REG=1 (A/B runs): leaves in registers for the per-frame launches of the shipped layer with a shared DB.  Measured equal to the
    // shared-memory form (1.33 vs 1.34 ms per step: what the shared-memory reads cost, the halved occupancy gives back), so it stays off
    static const bool reg_form = [] { const char *e = getenv("OPB_GATS_REG"); return e && e[0] == '1'; }();
    if (L == 8 && shipped && reg_form && leaf_bstride == 0 && B >= 8) OPB_CHECK_CUDA(launch(opb::k_gats_planes<8, 1, 1>));
    else if (L == 8 && shipped) OPB_CHECK_CUDA(launch(opb::k_gats_planes<8, 1>));
    else if (L == 16 && shipped) OPB_CHECK_CUDA(launch(opb::k_gats_planes<16, 1>));
    else if (L == 8) OPB_CHECK_CUDA(launch(opb::k_gats_planes<8, -1>));
    else OPB_CHECK_CUDA(launch(opb::k_gats_planes<0, -1>));
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
