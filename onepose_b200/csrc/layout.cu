// Layout conversion: public channel-major [B, C, N] (GATs_SuperGlue.py:209-218) -> point-major [B, N, C].
#include "common.cuh"
#include "tc_common.cuh"

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

// The same transposition straight into the tensor-core path's stream state: split-precision planes of 64 x [B, N, 256] and the
// per-32-row column sums [B][op_linear_tc_stat_parts(N)][256][2] (sum, 0) -- what op_channel_major_to_point_major + op_split_planes +
// op_colsum_partial produce, bit for bit (same split, same summation order), without the fp32 point-major copy.
// Block = 32 points x 256 channels: warps read channel rows (128 bytes along N), thread = channel writes 32 points.
__global__ void __launch_bounds__(256) k_cm_to_planes(const float *__restrict__ in, long long bstride_in, __half *__restrict__ hi,
                                                      __half *__restrict__ lo, float2 *__restrict__ colsum, int N, int parts,
                                                      const int *__restrict__ row_counts) {
    __shared__ float tile[256][33];
    const int b = blockIdx.y, part = blockIdx.x, n0 = part * 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float *src = in + (long long)b * bstride_in;
    // ragged batch: points past this frame's count are padding -- zero planes, no part in the sums
    const int nvalid = row_counts ? min(N, __ldg(row_counts + b)) : N;
    for (int c = warp; c < 256; c += 8) tile[c][lane] = (n0 + lane < nvalid) ? src[(long long)c * N + n0 + lane] : 0.f;
    __syncthreads();
    const int c = threadIdx.x, n1 = max(0, min(32, N - n0));      // a part past the end (N not a multiple of 128) holds zero sums
    float s = 0.f;
    for (int i = 0; i < n1; ++i) {
        const float v = tile[c][i];
        s += v;
        __half h, l;
        tc::split_f16(v * 64.f, h, l);
        const long long o = ((long long)b * N + n0 + i) * 256 + c;
        hi[o] = h; lo[o] = l;
    }
    colsum[((long long)b * parts + part) * 256 + c] = make_float2(s, 0.f);
}

}  // namespace opb

extern "C" int op_channel_major_to_planes_ragged(const float *in, long long bstride_in, void *hi, void *lo, float *colsum_partial, int B, int N,
                                                 const int *row_counts, void *stream) {
    OPB_CHECK_ARG(in && hi && lo && colsum_partial, "op_channel_major_to_planes: null pointer");
    OPB_CHECK_ARG(B >= 0 && N >= 0 && B <= 65535, "op_channel_major_to_planes: bad shape B=%d N=%d", B, N);
    if (B == 0 || N == 0) return OPB_OK;
    const int parts = opb::ceil_div(N, 128) * 4;           // = op_linear_tc_stat_parts(N)
    opb::k_cm_to_planes<<<dim3(parts, B), 256, 0, (cudaStream_t)stream>>>(in, bstride_in, (__half *)hi, (__half *)lo, (float2 *)colsum_partial, N, parts, row_counts);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_channel_major_to_planes(const float *in, long long bstride_in, void *hi, void *lo, float *colsum_partial, int B, int N,
                                          void *stream) {
    return op_channel_major_to_planes_ragged(in, bstride_in, hi, lo, colsum_partial, B, N, nullptr, stream);
}

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
