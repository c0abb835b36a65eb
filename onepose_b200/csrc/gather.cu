// Valid-match gather -- the 6 lines of numpy between the matcher and PnP in the reference
// (inference.py:147-152, eval_utils.py:72-77): valid = matches0 > -1; mkpts2d = kpts2d[valid];
// mkpts3d = kpts3d[matches0[valid]]; mconf = scores0[valid].  Ascending query order is kept by an
// in-block exclusive scan, so the compacted list is identical to numpy's boolean indexing.
#include "common.cuh"

namespace opb {

__global__ void __launch_bounds__(1024)
k_gather_matches(const int64_t *__restrict__ matches0, const float *__restrict__ scores0, const float *__restrict__ kpts2d,
                 const float *__restrict__ kpts3d, long long kp3_bstride, int N2, int N3, double scale,
                 double *__restrict__ out2d, double *__restrict__ out3d, float *__restrict__ out_conf,
                 int32_t *__restrict__ out_qidx, int32_t *__restrict__ count) {
    __shared__ int warp_tot[32];
    __shared__ int base;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) base = 0;
    __syncthreads();
    const int64_t *m = matches0 + (long long)b * N2;
    const float *k3 = kpts3d + (long long)b * kp3_bstride;
    for (int start = 0; start < N2; start += 1024) {
        const int i = start + tid;
        const long long j = i < N2 ? m[i] : -1;
        const bool valid = j > -1 && j < N3;
        const unsigned bal = __ballot_sync(0xffffffffu, valid);
        const int prefix = __popc(bal & ((1u << lane) - 1));
        if (lane == 0) warp_tot[warp] = __popc(bal);
        __syncthreads();
        int woff = 0, total = 0;
        for (int w = 0; w < 32; ++w) { const int t = warp_tot[w]; if (w < warp) woff += t; total += t; }
        const int pos = base + woff + prefix;
        if (valid) {
            const long long o = (long long)b * N2 + pos;
            out2d[o * 2] = (double)kpts2d[((long long)b * N2 + i) * 2];
            out2d[o * 2 + 1] = (double)kpts2d[((long long)b * N2 + i) * 2 + 1];
            out3d[o * 3] = (double)k3[j * 3] * scale;
            out3d[o * 3 + 1] = (double)k3[j * 3 + 1] * scale;
            out3d[o * 3 + 2] = (double)k3[j * 3 + 2] * scale;
            out_conf[o] = scores0[(long long)b * N2 + i];
            out_qidx[o] = i;
        }
        __syncthreads();
        if (tid == 0) base += total;
        __syncthreads();
    }
    if (tid == 0) count[b] = base;
    // rows past the compacted list are zero, so that callers can hand in uninitialised buffers
    for (int pos = base + tid; pos < N2; pos += 1024) {
        const long long o = (long long)b * N2 + pos;
        out2d[o * 2] = 0.0; out2d[o * 2 + 1] = 0.0;
        out3d[o * 3] = 0.0; out3d[o * 3 + 1] = 0.0; out3d[o * 3 + 2] = 0.0;
        out_conf[o] = 0.f; out_qidx[o] = 0;
    }
}

}  // namespace opb

extern "C" int op_gather_matches(const int64_t *matches0, const float *scores0, const float *kpts2d, const float *kpts3d,
                                 long long kp3_bstride, int B, int N2, int N3, double scale, double *out2d, double *out3d,
                                 float *out_conf, int32_t *out_query_idx, int32_t *count, void *stream) {
    OPB_CHECK_ARG(matches0 && scores0 && kpts2d && kpts3d && out2d && out3d && out_conf && out_query_idx && count,
                  "op_gather_matches: null pointer");
    OPB_CHECK_ARG(B > 0 && N2 > 0 && N3 > 0, "op_gather_matches: bad shape");
    opb::k_gather_matches<<<B, 1024, 0, (cudaStream_t)stream>>>(matches0, scores0, kpts2d, kpts3d, kp3_bstride, N2, N3, scale,
                                                               out2d, out3d, out_conf, out_query_idx, count);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
