// SURVEY.md 8(f) row f3 -- the SuperPoint front-end's post-network stages (superpoint.py:47-113,205-244): everything between the
// convolution stack (cuDNN-class library work, driven from the host module) and the matcher's inputs.
//
//  k_sp_softmax_shuffle : 65-way softmax over the channel axis, drop the dustbin, 8x8 depth-to-space  (:207-211)
//  k_sp_pool_stage      : one max-pool stage of simple_nms (:47-69) on a shared-memory tile with halo, with the element-wise
//                         logic before / after the pool fused in; five launches reproduce the reference's mask algebra exactly
//  k_sp_select          : threshold, border removal, row-major ordered compaction (torch.nonzero order), and -- when more than
//                         max_keypoints survive -- the top-k by score (bitonic sort of 64-bit (score, inverted index) keys)
//  k_sp_desc_prepare    : dense descriptors [256,h,w] -> L2-normalised pixel-major [h*w,256]  (:235-236)
//  k_sp_sample          : bilinear sampling at the keypoints (grid_sample, align_corners=True) + L2 normalisation (:96-113)
#include "common.cuh"

namespace opb {
namespace sp {

// cell_stride == 0: logits are [B,65,h,w] (channel-major, the library convolution's layout);
// cell_stride  > 0: logits are [B,h,w,cell_stride] (pixel-major, op_conv_tc's layout), 65 <= cell_stride
__global__ void k_sp_softmax_shuffle(const float *__restrict__ logits, int h, int w, int cell_stride, float *__restrict__ scores) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= w) return;
    const long long plane = (long long)h * w;
    const float *p = cell_stride ? logits + ((long long)b * plane + (long long)y * w + x) * cell_stride
                                 : logits + (long long)b * 65 * plane + (long long)y * w + x;
    const long long cs = cell_stride ? 1 : plane;
    float v[65];
    float m = -INFINITY;
#pragma unroll
    for (int c = 0; c < 65; ++c) { v[c] = p[c * cs]; m = fmaxf(m, v[c]); }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < 65; ++c) { v[c] = expf(v[c] - m); s += v[c]; }
    float *o = scores + (long long)b * plane * 64 + (long long)(y * 8) * (w * 8) + x * 8;
#pragma unroll
    for (int c = 0; c < 64; ++c) o[(c >> 3) * (w * 8) + (c & 7)] = v[c] / s;      // channel c -> cell offset (c / 8, c % 8)
}

// One (2r+1)^2 max-pool stage, -inf outside the image (max_pool2d's implicit padding).
//   MODE 0: mask = (scores == pool(scores))
//   MODE 1: supp = (pool(mask) > 0)
//   MODE 2: ss = supp ? 0 : scores;  new = (ss == pool(ss));  mask |= new & !supp;  if (last) out = mask ? scores : 0
constexpr int TILE = 32;
template <int MODE>
__global__ void __launch_bounds__(TILE * 8)
k_sp_pool_stage(const float *__restrict__ scores, unsigned char *__restrict__ mask, unsigned char *__restrict__ supp, int H, int W, int r,
                int last, float *__restrict__ out) {
    extern __shared__ float tile[];                       // (TILE+2r) x (TILE+2r) values, then (TILE+2r) x TILE row maxima
    const int S = TILE + 2 * r;
    float *rowmax = tile + S * S;
    const int b = blockIdx.z, x0 = blockIdx.x * TILE, y0 = blockIdx.y * TILE;
    const long long img = (long long)b * H * W;
    for (int i = threadIdx.x; i < S * S; i += blockDim.x) {
        const int ty = i / S, tx = i % S, y = y0 + ty - r, x = x0 + tx - r;
        float v = -INFINITY;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            const long long o = img + (long long)y * W + x;
            if (MODE == 0) v = scores[o];
            else if (MODE == 1) v = mask[o] ? 1.f : 0.f;
            else v = supp[o] ? 0.f : scores[o];
        }
        tile[i] = v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < S * TILE; i += blockDim.x) {       // horizontal pass: rows of the halo'd tile, TILE centre columns
        const int ty = i / TILE, tx = i % TILE;
        float m = -INFINITY;
        for (int d = 0; d <= 2 * r; ++d) m = fmaxf(m, tile[ty * S + tx + d]);
        rowmax[i] = m;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < TILE * TILE; i += blockDim.x) {    // vertical pass + the stage's element-wise logic
        const int ty = i / TILE, tx = i % TILE, y = y0 + ty, x = x0 + tx;
        if (y >= H || x >= W) continue;
        float m = -INFINITY;
        for (int d = 0; d <= 2 * r; ++d) m = fmaxf(m, rowmax[(ty + d) * TILE + tx]);
        const long long o = img + (long long)y * W + x;
        const float c = tile[(ty + r) * S + tx + r];
        if (MODE == 0) mask[o] = (c == m);
        else if (MODE == 1) supp[o] = (m > 0.f);
        else {
            const bool sp_ = supp[o] != 0;
            const bool mk = (mask[o] != 0) || ((c == m) && !sp_);
            mask[o] = mk;
            if (last) out[o] = mk ? scores[o] : 0.f;
        }
    }
}

// One block per image.  Candidates (score > thr, inside the border) are compacted in row-major order; if more than max_k survive
// they are sorted by (score desc, position asc) and the first max_k kept.  kpts are (x, y) floats as the reference's flip (:228).
constexpr int kSelThreads = 1024;
constexpr int kMaxCand = 16384;                       // NMS maxima are > r apart: <= ceil(H/(r+1)) * ceil(W/(r+1))
__global__ void __launch_bounds__(kSelThreads)
k_sp_select(const float *__restrict__ nms, int H, int W, float thr, int border, int max_k, int out_stride,
            float *__restrict__ kpts, float *__restrict__ kscores, int *__restrict__ count, int *__restrict__ overflow) {
    extern __shared__ unsigned long long keys[];          // kMaxCand 64-bit keys
    __shared__ int warp_tot[32];
    __shared__ int base;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *img = nms + (long long)b * H * W;
    if (tid == 0) base = 0;
    __syncthreads();
    const int n = H * W;
    for (int start = 0; start < n; start += kSelThreads) {
        const int i = start + tid;
        bool valid = false;
        float s = 0.f;
        if (i < n) {
            const int y = i / W, x = i % W;
            s = img[i];
            valid = s > thr && y >= border && y < H - border && x >= border && x < W - border;       // remove_borders, :72-83
        }
        const unsigned bal = __ballot_sync(0xffffffffu, valid);
        const int prefix = __popc(bal & ((1u << lane) - 1));
        if (lane == 0) warp_tot[warp] = __popc(bal);
        __syncthreads();
        int woff = 0, total = 0;
        for (int w2 = 0; w2 < 32; ++w2) { const int t = warp_tot[w2]; if (w2 < warp) woff += t; total += t; }
        const int pos = base + woff + prefix;
        if (valid && pos < kMaxCand)
            keys[pos] = ((unsigned long long)__float_as_uint(s) << 32) | (unsigned long long)(0xffffffffu - (unsigned)i);   // s > 0: bits are monotonic
        __syncthreads();
        if (tid == 0) base += total;
        __syncthreads();
    }
    int ncand = base;
    if (ncand > kMaxCand) { if (tid == 0) overflow[b] = ncand; ncand = kMaxCand; }
    const bool need_topk = max_k >= 0 && ncand > max_k;
    if (need_topk) {
        int np2 = 1;
        while (np2 < ncand) np2 <<= 1;
        for (int i = ncand + tid; i < np2; i += kSelThreads) keys[i] = 0ull;
        __syncthreads();
        for (int k = 2; k <= np2; k <<= 1)                 // bitonic sort, descending
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < np2; i += kSelThreads) {
                    const int l = i ^ j;
                    if (l > i) {
                        const unsigned long long a = keys[i], c = keys[l];
                        const bool desc = (i & k) == 0;
                        if (desc ? (a < c) : (a > c)) { keys[i] = c; keys[l] = a; }
                    }
                }
                __syncthreads();
            }
    }
    const int nout = need_topk ? max_k : ncand;
    for (int i = tid; i < nout; i += kSelThreads) {
        const unsigned long long k = keys[i];
        const unsigned idx = 0xffffffffu - (unsigned)(k & 0xffffffffull);
        kpts[((long long)b * out_stride + i) * 2] = (float)(idx % W);
        kpts[((long long)b * out_stride + i) * 2 + 1] = (float)(idx / W);
        kscores[(long long)b * out_stride + i] = __uint_as_float((unsigned)(k >> 32));
    }
    if (tid == 0) count[b] = nout;
}

// dense [B,256,h*w] -> pixel-major, L2-normalised per pixel (x / max(||x||, 1e-12))
__global__ void __launch_bounds__(256)
k_sp_desc_prepare(const float *__restrict__ dense, int hw, float *__restrict__ out) {
    __shared__ float t[32][257];
    const int b = blockIdx.y, p0 = blockIdx.x * 32, tid = threadIdx.x;
    const float *src = dense + (long long)b * kD * hw;
    for (int i = tid; i < 32 * kD; i += 256) {
        const int c = i / 32, p = i % 32;
        t[p][c] = (p0 + p < hw) ? src[(long long)c * hw + p0 + p] : 0.f;
    }
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    for (int p = warp; p < 32; p += 8) {
        if (p0 + p >= hw) continue;
        float v[8], ss = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) { v[q] = t[p][lane + 32 * q]; ss += v[q] * v[q]; }
        ss = warp_sum(ss);
        const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
#pragma unroll
        for (int q = 0; q < 8; ++q) out[((long long)b * hw + p0 + p) * kD + lane + 32 * q] = v[q] * inv;
    }
}

// warp per keypoint: grid_sample(bilinear, zeros padding, align_corners=True) of the normalised map, then L2 normalise
__global__ void __launch_bounds__(256)
k_sp_sample(const float *__restrict__ desc_pm, const float *__restrict__ kpts, const int *__restrict__ count, int stride, int h, int w, int s,
            int align_corners, float *__restrict__ out_pm) {
    const int b = blockIdx.y, k = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (k >= stride) return;
    float *o = out_pm + ((long long)b * stride + k) * kD;
    if (k >= count[b]) {                                  // padding rows of the batch tensor: zeros
        for (int q = 0; q < 8; ++q) o[lane + 32 * q] = 0.f;
        return;
    }
    // :100-104: keypoints - s/2 + 0.5, divided by (w*s - s/2 - 0.5, h*s - s/2 - 0.5), times 2 minus 1
    float gx = kpts[((long long)b * stride + k) * 2] - (float)s / 2 + 0.5f;
    float gy = kpts[((long long)b * stride + k) * 2 + 1] - (float)s / 2 + 0.5f;
    gx /= (float)(w * s) - (float)s / 2 - 0.5f; gy /= (float)(h * s) - (float)s / 2 - 0.5f;
    gx = gx * 2 - 1; gy = gy * 2 - 1;
    // grid_sample's unnormalisation.  align_corners=True is what the reference gets on its pinned torch 1.8 (:106 tests the
    // version string); False is what the same line selects on torch >= 2.10 ("2.1x"[2] == '1').
    const float ix = align_corners ? (gx + 1) / 2 * (w - 1) : ((gx + 1) * w - 1) / 2;
    const float iy = align_corners ? (gy + 1) / 2 * (h - 1) : ((gy + 1) * h - 1) / 2;
    const float fx = floorf(ix), fy = floorf(iy);
    const int x0 = (int)fx, y0 = (int)fy, x1 = x0 + 1, y1 = y0 + 1;
    const float wx1 = ix - fx, wy1 = iy - fy, wx0 = 1.f - wx1, wy0 = 1.f - wy1;
    const float wnw = wx0 * wy0, wne = wx1 * wy0, wsw = wx0 * wy1, wse = wx1 * wy1;
    const float *base = desc_pm + (long long)b * h * w * kD;
    auto at = [&](int y, int x, int c) -> float { return (y >= 0 && y < h && x >= 0 && x < w) ? base[((long long)y * w + x) * kD + c] : 0.f; };
    float v[8], ss = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const int c = lane + 32 * q;
        v[q] = at(y0, x0, c) * wnw + at(y0, x1, c) * wne + at(y1, x0, c) * wsw + at(y1, x1, c) * wse;
        ss += v[q] * v[q];
    }
    ss = warp_sum(ss);
    const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);
#pragma unroll
    for (int q = 0; q < 8; ++q) o[lane + 32 * q] = v[q] * inv;
}

}  // namespace sp
}  // namespace opb

extern "C" int op_sp_softmax_shuffle(const float *logits, int B, int h, int w, int cell_stride, float *scores, void *stream) {
    OPB_CHECK_ARG(logits && scores && B > 0 && h > 0 && w > 0 && h <= 65535 && B <= 65535, "op_sp_softmax_shuffle: bad argument");
    OPB_CHECK_ARG(cell_stride == 0 || cell_stride >= 65, "op_sp_softmax_shuffle: cell_stride=%d", cell_stride);
    opb::sp::k_sp_softmax_shuffle<<<dim3(opb::ceil_div(w, 64), h, B), 64, 0, (cudaStream_t)stream>>>(logits, h, w, cell_stride, scores);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" size_t op_sp_nms_workspace_bytes(int B, int H, int W) { return (size_t)B * H * W * 2; }

extern "C" int op_sp_nms(const float *scores, int B, int H, int W, int radius, float *out, void *workspace, size_t workspace_bytes, void *stream) {
    using namespace opb::sp;
    OPB_CHECK_ARG(scores && out && workspace && B > 0 && H > 0 && W > 0 && radius >= 0 && radius <= 16, "op_sp_nms: bad argument");
    OPB_CHECK_ARG(workspace_bytes >= op_sp_nms_workspace_bytes(B, H, W), "op_sp_nms: workspace too small");
    OPB_CHECK_ARG(B <= 65535 && opb::ceil_div(H, TILE) <= 65535, "op_sp_nms: image too large");
    unsigned char *mask = (unsigned char *)workspace, *supp = mask + (size_t)B * H * W;
    const int S = TILE + 2 * radius;
    const size_t smem = (size_t)(S * S + S * TILE) * sizeof(float);
    const dim3 grid(opb::ceil_div(W, TILE), opb::ceil_div(H, TILE), B);
    cudaStream_t st = (cudaStream_t)stream;
    k_sp_pool_stage<0><<<grid, TILE * 8, smem, st>>>(scores, mask, supp, H, W, radius, 0, out);
    for (int it = 0; it < 2; ++it) {                                                    // superpoint.py:62-67
        k_sp_pool_stage<1><<<grid, TILE * 8, smem, st>>>(scores, mask, supp, H, W, radius, 0, out);
        k_sp_pool_stage<2><<<grid, TILE * 8, smem, st>>>(scores, mask, supp, H, W, radius, it == 1, out);
    }
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_sp_select(const float *nms_scores, int B, int H, int W, float threshold, int border, int max_keypoints, int out_stride,
                            float *kpts, float *kscores, int *count, int *overflow, void *stream) {
    using namespace opb::sp;
    OPB_CHECK_ARG(nms_scores && kpts && kscores && count && overflow && B > 0 && H > 0 && W > 0, "op_sp_select: bad argument");
    OPB_CHECK_ARG(out_stride > 0 && (max_keypoints < 0 ? out_stride >= kMaxCand : out_stride >= max_keypoints),
                  "op_sp_select: out_stride=%d too small for max_keypoints=%d", out_stride, max_keypoints);
    OPB_CHECK_ARG(threshold >= 0.f, "op_sp_select: the keypoint threshold must be non-negative");
    static bool configured = false;
    if (!configured) {
        OPB_CHECK_CUDA(cudaFuncSetAttribute(k_sp_select, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxCand * 8));
        configured = true;
    }
    OPB_CHECK_CUDA(cudaMemsetAsync(overflow, 0, sizeof(int) * B, (cudaStream_t)stream));
    k_sp_select<<<B, kSelThreads, kMaxCand * 8, (cudaStream_t)stream>>>(nms_scores, H, W, threshold, border, max_keypoints, out_stride, kpts,
                                                                       kscores, count, overflow);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_sp_desc_prepare(const float *dense, int B, int hw, float *desc_pm, void *stream) {
    OPB_CHECK_ARG(dense && desc_pm && B > 0 && hw > 0 && B <= 65535, "op_sp_desc_prepare: bad argument");
    opb::sp::k_sp_desc_prepare<<<dim3(opb::ceil_div(hw, 32), B), 256, 0, (cudaStream_t)stream>>>(dense, hw, desc_pm);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}

extern "C" int op_sp_sample(const float *desc_pm, const float *kpts, const int *count, int B, int stride, int h, int w, int s, int align_corners,
                            float *out_pm, void *stream) {
    OPB_CHECK_ARG(desc_pm && kpts && count && out_pm && B > 0 && stride > 0 && h > 1 && w > 1 && s > 0 && B <= 65535, "op_sp_sample: bad argument");
    opb::sp::k_sp_sample<<<dim3(opb::ceil_div(stride, 8), B), 256, 0, (cudaStream_t)stream>>>(desc_pm, kpts, count, stride, h, w, s, align_corners, out_pm);
    OPB_CHECK_LAUNCH();
    return OPB_OK;
}
