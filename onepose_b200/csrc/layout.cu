// Layout conversion: public channel-major [B, C, N] (GATs_SuperGlue.py:209-218) -> point-major [B, N, C].
#include "common.cuh"

namespace opb {

// 32x32 smem tile transpose; reads coalesced along N, writes coalesced along C.
__global__ void k_cm_to_pm(const float *__restrict__ in, float *__restrict__ out, int C, int N,
                           long long bstride_in) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z;
    const int n0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const float *src = in + (long long)b * bstride_in;
    float *dst = out + (long long)b * C * N;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int c = c0 + i, n = n0 + threadIdx.x;
        if (c < C && n < N) tile[i][threadIdx.x] = src[(long long)c * N + n];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int n = n0 + i, c = c0 + threadIdx.x;
        if (c < C && n < N) dst[(long long)n * C + c] = tile[threadIdx.x][i];
    }
}

}  // namespace opb

extern "C" int op_channel_major_to_point_major(const float *in, float *out, int B, int C, int N,
                                               long long bstride_in, void *stream) {
    OPB_CHECK_ARG(in && out, "op_channel_major_to_point_major: null pointer");
    OPB_CHECK_ARG(B >= 0 && C > 0 && N >= 0, "op_channel_major_to_point_major: bad shape B=%d C=%d N=%d", B, C, N);
    if (B == 0 || N == 0) return OPB_OK;
    OPB_CHECK_ARG(B <= 65535 && opb::ceil_div(C, 32) <= 65535, "op_channel_major_to_point_major: B or C too large");
    dim3 grid(opb::ceil_div(N, 32), opb::ceil_div(C, 32), B), block(32, 8);
    opb::k_cm_to_pm<<<grid, block, 0, (cudaStream_t)stream>>>(in, out, C, N, bstride_in);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
