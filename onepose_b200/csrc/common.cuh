// Shared helpers for the onepose_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/onepose_b200.h"

namespace opb {

void set_error(const char *fmt, ...);

#define OPB_CHECK_ARG(cond, ...)            \
    do {                                    \
        if (!(cond)) {                      \
            opb::set_error(__VA_ARGS__);    \
            return OPB_ERR_ARG;             \
        }                                   \
    } while (0)

#define OPB_CHECK_CUDA(expr)                                                                   \
    do {                                                                                       \
        cudaError_t e_ = (expr);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            opb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return OPB_ERR_CUDA;                                                               \
        }                                                                                      \
    } while (0)

#define OPB_CHECK_LAUNCH() OPB_CHECK_CUDA(cudaGetLastError())

constexpr int kD = 256;       // descriptor_dim of the shipped model (train_GATsSPG.yaml:44)
constexpr int kHeads = 4;     // AttentionPropagation(feature_dim, 4), GATs_SuperGlue.py:63
constexpr int kDh = 64;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float elu1(float x) { return x > 0.f ? x + 1.f : expm1f(x) + 1.f; }   // elu(x)+1, same rounding as torch
__device__ __forceinline__ float elu(float x) { return x > 0.f ? x : expm1f(x); }

inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace opb
