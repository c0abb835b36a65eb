// SURVEY.md 8(f) rows f1 / f2 -- the data either side of the matcher.
//
//  k_gather_rows    : the leaf tensor of an object DB in the matcher's internal point-major layout, built from the collected
//                     leaf descriptors by the affiliation index plan of build_features3d_leaves (data_utils.py:163-205):
//                     out[r] = src[idx[r]], or the all-ones dustbin descriptor where idx[r] < 0 (dustbin pad / tail pad).
//  k_segment_mean   : mean_descriptors (feature_process.py:297-305): per 3D point the mean of its leaf descriptors, rows added
//                     in order in fp32 and divided once (numpy's np.mean(axis=0) order).
//  k_pose_errors    : the cm-degree yard-stick of Evaluator / query_pose_error (cmd_evaluator.py:11-33, eval_utils.py:45-63)
//                     for a batch of poses that never left the device.
#include "common.cuh"

namespace opb {

// one warp per output row of 256 floats
__global__ void __launch_bounds__(256)
k_gather_rows(const float *__restrict__ src, const int64_t *__restrict__ idx, long long n_src, long long n_out, float *__restrict__ out) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (r >= n_out) return;
    const long long s = idx[r];
    float4 a = make_float4(1.f, 1.f, 1.f, 1.f), b = a;
    if (s >= 0 && s < n_src) {
        const float4 *p = reinterpret_cast<const float4 *>(src + s * kD);
        a = __ldg(p + lane); b = __ldg(p + 32 + lane);
    }
    float4 *o = reinterpret_cast<float4 *>(out + r * kD);
    o[lane] = a; o[32 + lane] = b;
}

// numpy's pairwise summation of a contiguous fp32 vector (pairwise_sum_FLOAT: < 8 elements in order; up to 128 with eight
// interleaved accumulators; longer vectors split in halves) -- what np.mean(axis=0) does on an [n,1] score column
__device__ float np_pairwise_sum(const float *a, long long n) {
    if (n < 8) {
        float res = 0.f;
        for (long long i = 0; i < n; ++i) res += a[i];
        return res;
    }
    if (n <= 128) {
        float r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        long long i = 8;
        for (; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += a[i + j];
        float res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    long long n2 = n / 2;
    n2 -= n2 % 8;
    return np_pairwise_sum(a, n2) + np_pairwise_sum(a + n2, n - n2);
}

// one warp per segment [offsets[i], offsets[i+1]) of rows of C floats.  C > 1: rows are added in order per column (numpy's
// order for a reduction over the outer axis); C == 1: numpy's pairwise order (the column is contiguous).
__global__ void __launch_bounds__(256)
k_segment_mean(const float *__restrict__ src, const int64_t *__restrict__ offsets, long long n_seg, int C, float *__restrict__ out) {
    const long long i = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= n_seg) return;
    const long long r0 = offsets[i], r1 = offsets[i + 1];
    if (C == 1) {
        if (lane == 0) out[i] = np_pairwise_sum(src + r0, r1 - r0) / (float)(r1 - r0);
        return;
    }
    for (int c = lane; c < C; c += 32) {
        float s = 0.f;
        for (long long r = r0; r < r1; ++r) s += src[r * C + c];
        out[i * C + c] = s / (float)(r1 - r0);
    }
}

__global__ void k_pose_errors(const double *__restrict__ pred, const double *__restrict__ gt, int F, double *__restrict__ t_cm,
                              double *__restrict__ r_deg) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    const double *p = pred + (long long)f * 12, *g = gt + (long long)f * 12;      // [3,4] row-major: R | t
    const double dx = p[3] - g[3], dy = p[7] - g[7], dz = p[11] - g[11];
    t_cm[f] = sqrt(dx * dx + dy * dy + dz * dz) * 100.0;
    double trace = 0.0;                                                           // trace(R_pred R_gt^T) = sum_ij Rp_ij Rg_ij
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) trace += p[i * 4 + j] * g[i * 4 + j];
    trace = trace <= 3.0 ? trace : 3.0;                                           // only the upper clamp, as the reference
    r_deg[f] = acos((trace - 1.0) / 2.0) * (180.0 / 3.14159265358979323846);
}

}  // namespace opb

extern "C" int op_gather_rows(const float *src, const int64_t *idx, long long n_src, long long n_out, float *out, void *stream) {
    OPB_CHECK_ARG(idx && out && (src || n_src == 0) && n_src >= 0 && n_out >= 0, "op_gather_rows: bad argument");
    if (n_out == 0) return OPB_OK;
    opb::k_gather_rows<<<(unsigned)opb::ceil_div(n_out, 8), 256, 0, (cudaStream_t)stream>>>(src, idx, n_src, n_out, out);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_segment_mean(const float *src, const int64_t *offsets, long long n_seg, int C, float *out, void *stream) {
    OPB_CHECK_ARG(src && offsets && out && n_seg >= 0 && C > 0, "op_segment_mean: bad argument");
    if (n_seg == 0) return OPB_OK;
    opb::k_segment_mean<<<(unsigned)opb::ceil_div(n_seg, 8), 256, 0, (cudaStream_t)stream>>>(src, offsets, n_seg, C, out);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_pose_errors(const double *pose_pred, const double *pose_gt, int F, double *t_err_cm, double *r_err_deg, void *stream) {
    OPB_CHECK_ARG(pose_pred && pose_gt && t_err_cm && r_err_deg && F >= 0, "op_pose_errors: bad argument");
    if (F == 0) return OPB_OK;
    opb::k_pose_errors<<<opb::ceil_div(F, 128), 128, 0, (cudaStream_t)stream>>>(pose_pred, pose_gt, F, t_err_cm, r_err_deg);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
