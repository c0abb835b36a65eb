/*
 * onepose_b200 -- C ABI of the B200 (sm_100a) GATsSPG matching + PnP hot path.
 *
 * The reference (huanghaoran111/OnePose) is pure Python and has no FFI of its own; the functions
 * below are the operator boundary a maintainer binds (ctypes stub in INTEGRATION.md) to replace
 *   - GATsSuperGlue.forward      src/models/GATsSPG_architectures/GATs_SuperGlue.py:203-278
 *   - GraphAttentionLayer.forward src/models/GATsSPG_architectures/GATs.py:62-123
 *   - ransac_PnP                 src/utils/eval_utils.py:18-42
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host; the caller owns all memory
 *     (including workspaces); nothing here allocates, frees or synchronises the device;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*);
 *   - return value 0 = success, negative = error; op_last_error() gives the message of the last
 *     failure on the calling thread; nothing throws or exits across the ABI;
 *   - activations are fp32 "point-major" [B, N, C] (row = one keypoint), indices int64 as in the
 *     reference (-1 = unmatched), poses float64.
 */
#ifndef ONEPOSE_B200_H
#define ONEPOSE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OPB_OK 0
#define OPB_ERR_ARG (-1)
#define OPB_ERR_CUDA (-2)
#define OPB_ERR_WORKSPACE (-3)

const char *op_last_error(void);
int op_abi_version(void);

/* ---- layout --------------------------------------------------------------------------------
 * Public tensors are channel-major [B, C, N] (GATs_SuperGlue.py:209-218).  Convert to the internal
 * point-major [B, N, C].  bstride_in (in elements) may be 0 for an expand()-ed static DB. */
int op_channel_major_to_point_major(const float *in, float *out, int B, int C, int N,
                                    long long bstride_in, void *stream);
/* The same transposition straight into split-precision planes (64 x, [B, N, 256] fp16 hi / lo) plus the per-32-row column sums
 * [B][op_linear_tc_stat_parts(N)][256][2] of the tensor-core path: op_channel_major_to_point_major + op_split_planes + op_colsum_partial
 * in one pass, bit-identical, C = 256. */
int op_channel_major_to_planes(const float *in, long long bstride_in, void *hi, void *lo, float *colsum_partial, int B, int N,
                               void *stream);
/* Ragged batches.  The reference runs one frame per call with whatever keypoint count SuperPoint found (inference.py:140-146,
 * GATs_SuperGlue.py:203-278 at B = 1); a batch of frames with different counts is handed over padded to a common N, with
 * row_counts [B] (device, int32) giving each frame's valid rows.  The *_ragged forms of the entry points on the query stream take
 * that array (NULL = every row valid, the plain entry point): rows past a frame's count are left out of every cross-row quantity
 * (column sums, K'^T X and Ksum, InstanceNorm1d statistics, the dual-softmax sums and arg-maxima), so that frame b's results equal
 * the reference run on its first row_counts[b] points alone.  Padding rows of the outputs: zero planes, conf rows of +0,
 * matches0 = -1.  This one zero-fills the padding points' planes. */
int op_channel_major_to_planes_ragged(const float *in, long long bstride_in, void *hi, void *lo, float *colsum_partial, int B, int N,
                                      const int *row_counts, void *stream);

/* ---- (1) GATs leaf aggregation: GATs.py:62-123 ------------------------------------------------
 * For every 3D point n (and frame b): logits over its (include_self ? 1 : 0) + L entries,
 *   t = h3L[b,n].v2 ; s_0 = t (the self entry re-uses the a[256:] projection, GATs.py:115-120) ;
 *   s_j = leaf[n,j].v1 ; e = LeakyReLU_alpha(t + s)
 *   att = softmax(e) ; out = att_0*h3L[b,n] + sum_j att_j*leaf[n,j]      (include_self)
 *                      out = sum_j att_j*leaf[n,j] / 2 + h3L[b,n]         (!include_self)
 *   if (h3_add) out += h3_add[b,n] ; if (apply_elu) out = ELU(out)
 * leaf: [Bl, N3, L, 256] with batch stride leaf_bstride elements (0 = one DB shared by all frames).
 * h3L and out: [B, N3, 256]; out may alias h3L (each point reads only its own row).
 * With the shipped flags v1 = W a[:256], v2 = W a[256:] (the 256x256 GEMM collapses, GATs.py:68-69,113-115). */
int op_gats_leaf_aggregate(const float *leaf, long long leaf_bstride, const float *h3L, const float *h3_add,
                           const float *v1, const float *v2, float *out, int B, int N3, int L,
                           float alpha, int include_self, int apply_elu, void *stream);

/* The same layer in place on the split-precision planes (xh, xl: [B, N3, 256] fp16, section 2b) of the 3D stream, for
 * h3L == the stream itself (with_linear_transform = False); add_self = (additional && include_self); ELU always applied.
 * Also writes the per-frame column sums of the output as partials colsum_partial[B][op_gats_planes_parts(N3)][256][2]
 * ((sum, 0), the layout op_mlp_center_bias folds). */
int op_gats_planes_parts(int N3);
int op_gats_leaf_aggregate_planes(const float *leaf, long long leaf_bstride, void *xh, void *xl, const float *v1, const float *v2,
                                  float *colsum_partial, int B, int N3, int L, float alpha, int include_self, int add_self,
                                  void *stream);

/* ---- (2) self/cross AttentionPropagation building blocks: GATs_SuperGlue.py:88-147 -----------
 * Generic fused linear: C[b,n,:] = epi( [A1[b,n,:K1] | A2[b,n,:K2]] . W^T + bias ), W is [Cout, K1+K2]
 * row-major.  Rows are addressed as b*rows_per_batch + n.  Flags select the fusions:
 *   OPB_LIN_ELU1      : elu(x)+1 on output columns < elu_cols          (linear attention feature map, :90-91)
 *   OPB_LIN_RESIDUAL  : C += result (residual add, :78,83)
 *   OPB_LIN_STATS     : per (b, column) sum / sum of squares of the result in float64 partials
 *                       stats_partial[b][row_tile][Cout][2] (InstanceNorm1d statistics, :144)
 *   OPB_LIN_NORM_RELU : A1 is read as relu((a - mean[b,k]) * rstd[b,k])  (InstanceNorm + ReLU, :144-146)
 */
#define OPB_LIN_ELU1 1
#define OPB_LIN_RESIDUAL 2
#define OPB_LIN_STATS 4
#define OPB_LIN_NORM_RELU 8
#define OPB_LIN_COLSUM 16    /* op_linear_tc: stats_partial = (sum, 0) of the stored value over the valid rows */
#define OPB_LIN_ZSCALE 32    /* op_linear_tc: each 64-column head of a row is multiplied by zscale_num / (row_head . ksum[b, head] + 1e-6) */
#define OPB_LIN_WBATCHED 64  /* op_linear_tc: W planes are [B, Cout, K], one matrix per frame */
#define OPB_LIN_BIASB 128     /* op_linear_tc: bias is [B, Cout], one vector per frame */
#define OPB_LIN_CSTATS_RELU 256 /* op_linear_tc: stats_partial = (sum, sum of squares) of the value after bias over the valid rows, then ReLU */
#define OPB_LIN_W2BATCHED 512 /* op_linear_tc: the second weight W2 is [B, Cout, K2], one matrix per frame */
#define OPB_LIN_A1_BCAST 2048 /* op_linear_tc: A1 is [1, rows, K1], the same rows for every frame (e.g. the shared 3D stream at its first per-frame layer) */
#define OPB_LIN_A2_BCAST 4096 /* op_linear_tc: the same for A2 */
#define OPB_LIN_W2_IDENTITY 8192 /* op_linear_tc: W2 is the 256 x 256 identity (planes (64, 0); K2 = Cout = 256): C = epi(A1 . W^T + A2).  The pair form then issues the A2 part as N = 64 MMAs on the diagonal blocks only (a quarter of the tensor work of the residual) */
#define OPB_LIN_UNIT_PLANES8 16384 /* op_linear_tc (Cout = 256, planes output only): every output row is L2-normalised (x / max(|x|, 1e-12), F.normalize) and written as the operand planes of op_sim_dual_softmax_planes: Ch = fp16(64 u) [B,rows,256], Cl = bytes [2][B,rows,256]: (64 u - hi) as e5m2, then hi as e4m3 */
#define OPB_LIN_W2_EXACT_HI 1024 /* op_linear_tc: W2's lo plane is all zeros (e.g. the identity of the in-GEMM residual add): its MMA pass is skipped */
int op_linear(const float *A1, int K1, const float *A2, int K2, const float *W, const float *bias,
              float *C, int B, int rows_per_batch, int Cout, int flags, int elu_cols,
              const float *mean, const float *rstd, double *stats_partial, void *stream);
/* number of row tiles op_linear uses per batch element (sizes stats_partial) */
int op_linear_row_tiles(int rows_per_batch);
/* mean / rstd [B, C] from the float64 partials; biased variance, eps as InstanceNorm1d */
int op_instnorm_finalize(const double *stats_partial, int B, int row_tiles, int C, int rows_per_batch,
                         float eps, float *mean, float *rstd, void *stream);

/* KV reduction of linear attention (:93-97).  kv: [B, M, 512] = [K' (elu+1 applied) | V], channels
 * head-major (c = h*64 + d).  Outputs KV[b,h,d,q] = sum_m K'[m,h,d] * V[m,h,q] / M and
 * Ksum[b,h,d] = sum_m K'[m,h,d].  workspace: op_kv_reduce_workspace_bytes(B, M). */
size_t op_kv_reduce_workspace_bytes(int B, int M);
int op_kv_reduce(const float *kv, float *KV, float *Ksum, int B, int M, void *workspace,
                 size_t workspace_bytes, void *stream);
/* op_kv_reduce without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernels, same summation order and therefore the same bits as op_kv_reduce; the chunk partials live in a file-scope
 * static __device__ array of OPB_KV_STATIC_SCRATCH_BYTES (8 MiB, held by every GPU that loads the library until the process exits).
 * Limit: op_kv_reduce_workspace_bytes(B, M) <= OPB_KV_STATIC_SCRATCH_BYTES, i.e. B * ceil(M / 256) <= 126 (three frames at M = 8192);
 * larger shapes return OPB_ERR_WORKSPACE with a message and launch nothing.  Every call on a device shares the array: calls in one
 * stream are ordered, calls that overlap on the GPU (two streams, unordered graph branches) RACE on it.  Nothing has to be cleared
 * between calls (every partial that is read is written first).  Calls no cudaMalloc / cudaFree / allocating routine. */
#define OPB_KV_STATIC_SCRATCH_BYTES (8u << 20)
size_t op_kv_reduce_static_max_bytes(void);
int op_kv_reduce_static(const float *kv, float *KV, float *Ksum, int B, int M, void *stream);
/* out[b,n,h*64+q] = (sum_d Q'[b,n,h,d] KV[b,h,d,q]) * M / (sum_d Q'[b,n,h,d] Ksum[b,h,d] + 1e-6)  (:97-98) */
int op_attn_apply(const float *Qp, const float *KV, const float *Ksum, float *out, int B, int N, int M,
                  void *stream);

/* ---- (2b) the same projections on the 5th-gen tensor cores (tcgen05 + TMA) -------------------------
 * Operands are "split-precision planes": a tensor x is held as two fp16 arrays hi = fp16(64 x),
 * lo = fp16(64 x - hi) (22 significant bits; |x| < 1023).  A product is three kind::f16 MMAs
 * (hi*hi + lo*hi + hi*lo) accumulated in fp32 TMEM, which keeps conf_matrix within the 1e-3 bound
 * where a single fp16/bf16/tf32 pass does not (precision).  Plane pointers are void*
 * (fp16 storage), 16-byte aligned; K1, K2 multiples of 64; Cout 256 or 512.
 * op_linear_tc: C = epi([A1|A2] . W^T), or with W2 planes given C = epi(A1 . W^T + A2 . W2^T) (W [Cout,K1], W2 [Cout,K2] or
 * [B,Cout,K2] with OPB_LIN_W2BATCHED); outputs fp32 C and/or planes (Ch, Cl) for the next GEMM.
 * flags: OPB_LIN_ELU1, OPB_LIN_RESIDUAL (C += result; without C the old value is read from the output planes, in place), OPB_LIN_STATS (stats_partial
 * [B][op_linear_tc_stat_parts(rows)][Cout][2] fp32 partial sum / sum of squares of the result
 * WITHOUT the bias; op_instnorm_finalize_tc folds them in float64 and adds the bias to the mean). */
int op_split_planes(const float *x, void *hi, void *lo, long long n, void *stream);
int op_norm_relu_split(const float *h, const float *mean, const float *rstd, void *hi, void *lo, int B, int rows,
                       int C, void *stream);
int op_linear_tc(const void *A1h, const void *A1l, int K1, const void *A2h, const void *A2l, int K2, const void *Wh,
                 const void *Wl, const void *W2h, const void *W2l, const float *bias, float *C, void *Ch, void *Cl, int B,
                 int rows_per_batch, int Cout, int flags, int elu_cols, float *stats_partial, const float *ksum,
                 float zscale_num, void *stream);
/* Ragged form: OPB_LIN_COLSUM / OPB_LIN_CSTATS_RELU statistics over rows < row_counts[b] only, and the COLSUM producers (K', Q'',
 * the new x) write zeros to the padding rows.  OPB_LIN_STATS is refused. */
int op_linear_tc_ragged(const void *A1h, const void *A1l, int K1, const void *A2h, const void *A2l, int K2, const void *Wh,
                        const void *Wl, const void *W2h, const void *W2l, const float *bias, float *C, void *Ch, void *Cl, int B,
                        int rows_per_batch, int Cout, int flags, int elu_cols, float *stats_partial, const float *ksum,
                        float zscale_num, const int *row_counts, void *stream);
int op_linear_tc_stat_parts(int rows_per_batch);
/* The merge projection folded into MLP layer 0 (:119,135-147): message = Q'' P_b^T + bm and h = W0 [x | message] + b0, so
 *   h = W0a x + (W0b P_b) Q'' + const = W0a x + P'_b Q'' + const, with P'_b[c,(h,d)] = sum_q KV_h[d,q] (W0b Wm)[c,(h,q)] straight from
 *   op_kv_post (n_out = 512, WmT = the static product W0b Wm laid out [4,64,512]); the constant (W0b bm + b0) is removed by
 *   InstanceNorm1d and never added.  Layer 0 is then op_linear_tc(A1 = x, W = W0a, A2 = Q'', W2 = P'_b, OPB_LIN_W2BATCHED).
 * InstanceNorm1d of the MLP without a normalisation pass over the hidden tensor:
 *   - the per-channel mean of layer 0's output is analytic, W0a mean_rows(x) + P'_b mean_rows(Q'') + const, from the column sums of
 *     its inputs (OPB_LIN_COLSUM partials of their producers, or op_colsum_partial [B][op_linear_tc_stat_parts(rows)][C][2] for a
 *     plain tensor): op_mlp_center_bias writes the per-frame bias -(W0a mean_x + P'_b mean_q) [B,512] (W0aT = W0[:, :256]
 *     transposed, [256 k][512 c]; P planes [B or 1][512][256], p_batched says which);
 *   - layer 0 then runs with OPB_LIN_BIASB | OPB_LIN_CSTATS_RELU: output planes hold relu(h - mean), stats the centred sums;
 *   - op_instnorm_finalize_tc gives rstd, op_scale_weight_planes folds it into layer 1's weights: W1s[b,o,c] = W1[o,c] rstd[b,c]
 *     (planes [B,Cout,K]), used with OPB_LIN_WBATCHED. */
int op_colsum_partial(const float *x, int B, int rows, int C, float *partial, void *stream);
int op_mlp_center_bias(const float *xpart, int xparts, const float *qpart, int qparts, const float *W0aT, const void *Ph,
                       const void *Pl, int p_batched, int B, int rows, float *bias_out, void *stream);
/* ragged form: the means divide by row_counts[b] (at least 1) instead of rows */
int op_mlp_center_bias_ragged(const float *xpart, int xparts, const float *qpart, int qparts, const float *W0aT, const void *Ph,
                              const void *Pl, int p_batched, int B, int rows, const int *row_counts, float *bias_out, void *stream);
int op_scale_weight_planes(const float *W, const float *rstd, int B, int Cout, int K, void *hi, void *lo, void *stream);
/* op_instnorm_finalize_tc + op_scale_weight_planes in one launch: rstd from the CSTATS_RELU partials [B][parts][K][2] straight
 * into the per-frame layer-1 weight planes [B,Cout,K] (K a multiple of 64); rstd [B,K] is written too when not NULL. */
int op_instnorm_scale_weight_planes(const float *stats_partial, int B, int parts, int rows_per_batch, float eps, const float *W,
                                    int Cout, int K, void *hi, void *lo, float *rstd, void *stream);
/* ragged form: variance over row_counts[b] rows (at least 1) */
int op_instnorm_scale_weight_planes_ragged(const float *stats_partial, int B, int parts, int rows_per_batch, const int *row_counts,
                                           float eps, const float *W, int Cout, int K, void *hi, void *lo, float *rstd, void *stream);
int op_instnorm_finalize_tc(const float *stats_partial, const float *bias, int B, int parts, int C, int rows_per_batch,
                            float eps, float *mean, float *rstd, void *stream);
/* Source side of linear attention on the tensor cores (:88-99), re-associated so that the V projection is never
 * materialised: op_atb_tc computes per-chunk partials of G = K'^T X over the M source points (A = K' planes, B = X
 * planes, both [B, M, 256]; partial [B][op_atb_chunks(M)][256][256] fp32); op_kv_post folds them and the OPB_LIN_COLSUM
 * partials of K' (ksum_partial [B][nparts][256][2]) and emits ksum [B,256], optionally KV [B,4,64,64] (:96, scaled by inv_m = 1/M),
 * and the per-frame matrix P_b[o,(h,d)] = sum_q KV_h[d,q] Wm[o,(h,q)] as planes [B,n_out,256] (n_out = 256 or 512 rows of Wm), so that
 * message = (Q' M / (Q'_h.Ksum_h + 1e-6)) P_b^T + bm is one op_linear_tc call (OPB_LIN_ZSCALE on the Q projection,
 * OPB_LIN_WBATCHED on the merge).  Head-major weights, transposed for coalesced reads: WvT [4,256,64] = Wv[h*64+q, c], bv [256], WmT [4,64,n_out] = Wm[o, h*64+q]. */
int op_atb_chunks(int M);
int op_atb_tc(const void *Ah, const void *Al, const void *Bh, const void *Bl, int B, int M, float *partial, void *stream);
int op_kv_post(const float *partial, int nchunks, const float *ksum_partial, int nparts, const float *WvT, const float *bv,
               const float *WmT, int n_out, int B, float inv_m, float *ksum, float *kv_out, void *Ph, void *Pl, void *stream);
/* op_attn_apply writing split-precision planes instead of fp32 */
int op_attn_apply_planes(const float *Qp, const float *KV, const float *Ksum, void *out_hi, void *out_lo, int B, int N,
                         int M, void *stream);

/* row-wise L2 normalisation x / max(||x||, 1e-12)  (F.normalize, :245-246); in place allowed */
int op_l2_normalize_rows(const float *x, float *out, long long rows, int C, void *stream);

/* ---- (3) similarity + dual-softmax + mutual-NN: GATs_SuperGlue.py:251-274 -----------------------
 * m2 [B,N2,256], m3 [B,N3,256] unit rows.  conf[b,n,m] = softmax_n(s)*softmax_m(s), s = m2.m3/scale.
 * Pass 1 accumulates the row/column exp sums, pass 2 recomputes the tile and writes conf exactly
 * once, tracking row/column maxima; the selection kernel applies the mutual / threshold logic.
 * Outputs are batched: matches0 [B,N2] int64, matches1 [B,N3] int64, scores0 [B,N2], scores1 [B,N3].
 * Tensor-core form: CTA pairs (cta_group::2) on 256 x 256 tiles; a product is one fp16 MMA (hi*hi) plus the two cross terms as fp8
 * MMAs (lo e5m2 x hi e4m3): conf within 3e-4 of float64 (bound 1e-3).  Environment switches for A/B runs only: OPB_SIM_MMA=f16
 * (three fp16 MMAs per product, conf within 6e-5), OPB_SIM_CG=1 (one CTA per 128 x 128 tile). */
size_t op_sim_dual_softmax_workspace_bytes(int B, int N2, int N3);
int op_sim_dual_softmax(const float *m2, const float *m3, int B, int N2, int N3, float scale_factor,
                        float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                        float *scores0, float *scores1, void *workspace, size_t workspace_bytes,
                        void *stream);
/* Same contract for rows that are not yet unit length: F.normalize (:245-246) is fused into the operand split, so the
 * normalised fp32 copy of the final projections is never written (bit-identical to op_l2_normalize_rows + op_sim_dual_softmax). */
int op_sim_dual_softmax_rows(const float *x2, const float *x3, int B, int N2, int N3, float scale_factor,
                             float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                             float *scores0, float *scores1, void *workspace, size_t workspace_bytes,
                             void *stream);
/* Same contract on operand planes a producer has already written (op_linear_tc with OPB_LIN_UNIT_PLANES8: final_proj, :241, with
 * F.normalize, :245-246, in its epilogue): mX_hi = fp16(64 u) [B,N,256], mX_bytes = [2][B,N,256] bytes: (64 u - hi) as e5m2, then hi as
 * e4m3.  Neither the fp32 projections nor their normalised copy ever exist in HBM. */
int op_sim_dual_softmax_planes(const void *m2_hi, const void *m2_bytes, const void *m3_hi, const void *m3_bytes, int B, int N2, int N3,
                               float scale_factor, float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                               float *scores0, float *scores1, void *workspace, size_t workspace_bytes, void *stream);
/* Ragged form (see op_channel_major_to_planes_ragged): n2_counts [B] valid query rows per frame.  Rows past the count take no part in
 * the column softmax / column arg-maxima, their conf rows are +0 and their matches0 -1; frame b equals the reference run on its first
 * n2_counts[b] query points (GATs_SuperGlue.py:245-274 at B = 1).  Sixteen-warp pair form only. */
int op_sim_dual_softmax_planes_ragged(const void *m2_hi, const void *m2_bytes, const void *m3_hi, const void *m3_bytes, int B, int N2, int N3,
                                      const int *n2_counts, float scale_factor, float match_threshold, float *conf, int64_t *matches0,
                                      int64_t *matches1, float *scores0, float *scores1, void *workspace, size_t workspace_bytes,
                                      void *stream);
/* Same contract on fp32 CUDA cores (exact fp32 dot products, no tensor cores): kept as the on-device
 * cross-check of the tcgen05 split-precision kernel, not used by the product path. */
int op_sim_dual_softmax_simt(const float *m2, const float *m3, int B, int N2, int N3, float scale_factor,
                             float match_threshold, float *conf, int64_t *matches0, int64_t *matches1,
                             float *scores0, float *scores1, void *workspace, size_t workspace_bytes,
                             void *stream);

/* valid-match gather (inference.py:147-152): per frame, ascending query index.
 * pts2d [B,N2,2] fp32, pts3d [Bk,N3,3] fp32 (kp3_bstride elements, 0 = shared), out2d [B,N2,2] / out3d [B,N2,3] float64
 * (out3d scaled by `scale`), out_conf [B,N2], out_query_idx [B,N2] int32, count [B] int32. */
int op_gather_matches(const int64_t *matches0, const float *scores0, const float *kpts2d, const float *kpts3d,
                      long long kp3_bstride, int B, int N2, int N3, double scale, double *out2d, double *out3d,
                      float *out_conf, int32_t *out_query_idx, int32_t *count, void *stream);

/* ---- (3b) SURVEY.md 8(f): the data either side of the matcher -----------------------------------
 * op_gather_rows: out[r,:] = src[idx[r],:] (rows of 256 fp32), or the all-ones dustbin descriptor where idx[r] < 0 -- the
 *   leaf tensor of build_features3d_leaves (data_utils.py:163-205) directly in the matcher's point-major layout.
 * op_segment_mean: out[i,:] = mean(src[offsets[i]:offsets[i+1],:]) in fp32, rows added in order
 *   (mean_descriptors / mean_scores, feature_process.py:297-319).
 * op_pose_errors: per frame translation error in cm and rotation error in degrees of [3,4] float64 poses, as
 *   Evaluator.cm_degree_*_metric / query_pose_error (cmd_evaluator.py:11-33, eval_utils.py:45-63). */
int op_gather_rows(const float *src, const int64_t *idx, long long n_src, long long n_out, float *out, void *stream);
int op_segment_mean(const float *src, const int64_t *offsets, long long n_seg, int C, float *out, void *stream);
int op_pose_errors(const double *pose_pred, const double *pose_gt, int F, double *t_err_cm, double *r_err_deg, void *stream);

/* ---- (3c) SURVEY.md 8(f) row f3: SuperPoint's post-network stages (src/models/extractors/SuperPoint/superpoint.py) ---------
 * op_sp_softmax_shuffle: logits [B,65,h,w] (cell_stride = 0) or [B,h,w,cell_stride] (pixel-major, cell_stride >= 65) -> scores
 *   [B,8h,8w]: channel softmax, dustbin dropped, 8x8 depth-to-space (:207-211).
 * op_sp_nms: simple_nms (:47-69), five fused max-pool stages; workspace = op_sp_nms_workspace_bytes(B,H,W) bytes.
 * op_sp_select: score > threshold (torch.nonzero, row-major order), remove_borders (:72-83), top_k_keypoints (:86-93; sorted by
 *   score descending, ties by position) and the (h,w)->(x,y) flip (:228).  kpts [B,out_stride,2] fp32, kscores [B,out_stride],
 *   count [B] int32; overflow[b] != 0 reports an image with more than 16384 NMS survivors (not expected: maxima are > r apart).
 *   out_stride >= max_keypoints (>= 16384 when max_keypoints < 0).
 * op_sp_desc_prepare: dense descriptors [B,256,h*w] -> L2-normalised pixel-major [B,h*w,256] (:235-236).
 * op_sp_sample: sample_descriptors (:96-113) at kpts -> [B,out_stride,256] point-major, L2-normalised; rows >= count[b] are 0. */
int op_sp_softmax_shuffle(const float *logits, int B, int h, int w, int cell_stride, float *scores, void *stream);
size_t op_sp_nms_workspace_bytes(int B, int H, int W);
int op_sp_nms(const float *scores, int B, int H, int W, int radius, float *out, void *workspace, size_t workspace_bytes, void *stream);
int op_sp_select(const float *nms_scores, int B, int H, int W, float threshold, int border, int max_keypoints, int out_stride,
                 float *kpts, float *kscores, int32_t *count, int32_t *overflow, void *stream);
int op_sp_desc_prepare(const float *dense, int B, int hw, float *desc_pm, void *stream);
int op_sp_sample(const float *desc_pm, const float *kpts, const int32_t *count, int B, int stride, int h, int w, int s,
                 int align_corners, float *out_pm, void *stream);

/* ---- (3d) row f3: SuperPoint's convolution stack (superpoint.py:147-166,177-206,231-234) on the tensor cores ------------------
 * Activations are NHWC split-precision planes (hi, lo) of 64*x, as the matcher's streams (op_split_planes).
 * op_conv_tc: 3x3/pad 1 (ntaps = 9) or 1x1 (ntaps = 1) convolution + bias (+ ReLU).  x planes [B,H,W,Cin] (Cin % 64 == 0);
 *   weight planes [ntaps, Cout_pad, Cin] = 64 * weight[co, ci, ky, kx] at tap ky*3+kx, rows >= Cout zero, Cout_pad = 64 or a multiple
 *   of 128 (<= 256); bias [Cout] fp32.  Outputs: planes yh/yl [B,H,W,Cout] (Cout % 32 == 0) and/or fp32 y32 [B,H,W,y32_stride]
 *   (columns >= Cout of a row receive zeros up to the stride).  Three fp16 MMAs per product, fp32 accumulation in TMEM.
 *   pool = 1 (3x3, plane output, W % 8 == 0, even H): MaxPool2d(2,2) is applied in the epilogue, yh/yl are [B,H/2,W/2,Cout].
 * op_conv1a_planes: the first layer, image [B,1,H,W] fp32 -> planes [B,H,W,64], weights [64,1,3,3], ReLU (fp32 CUDA cores, K = 9).
 * op_pool2_planes: MaxPool2d(2,2) on planes [B,H,W,C] -> [B,H/2,W/2,C]. */
int op_conv_tc(const void *xh, const void *xl, int B, int H, int W, int Cin, const void *wh, const void *wl, int Cout_pad,
               const float *bias, int Cout, int ntaps, int relu, int pool, void *yh, void *yl, float *y32, int y32_stride, void *stream);
int op_conv1a_planes(const float *img, const float *w, const float *bias, int B, int H, int W, void *yh, void *yl, void *stream);
int op_pool2_planes(const void *xh, const void *xl, int B, int H, int W, int C, void *yh, void *yl, void *stream);

/* ---- (3e) SURVEY.md 8(f) row f4: tracker-side geometry (src/tracker/ba_tracker.py, tracking_utils.py), float64 ----------------
 * op_snavely_residual: SnavelyReprojectionErrorV2 + AngleAxisRotatePoint (tracking_utils.py:91-169).  points [.,3], cams [.,6]
 *   (angle-axis, translation), features [n,5] = (x, y, focal, l1, l2); observation i uses points[pt_idx[i]] / cams[cam_idx[i]]
 *   (NULL index = i, the reference's already-gathered form); residual [n,2].
 * op_project_points: `project` (tracking_utils.py:74-89): xy [n,2] (+ depth [n]); with ref2d [n,2], dist_sum receives the sum of
 *   |xy - ref2d| (flow_track's kpt_dist is dist_sum / n, ba_tracker.py:318-319).  K [3,3], RT [3,4] row-major, device memory.
 * op_triangulate_points: apply_triangulation (ba_tracker.py:267-273, cv2.triangulatePoints; DLT rows as :255-259): P1, P2 [3,4]
 *   projection matrices, x1, x2 [n,2] -> X [n,3] de-homogenised.
 * op_ba_lm: apply_ba (ba_tracker.py:358-441) -- windowed bundle adjustment on the Snavely residual.  The reference calls DeepLM's
 *   Solve (third-party, absent); this is Levenberg-Marquardt with analytic Jacobians, Schur complement on the points, dense Cholesky
 *   of the reduced camera system, accept / reject and damping (x0.1 / x10, Marquardt diagonal scaling) decided on the device.
 *   cams [C,6] and points [P,3] are updated in place; cam_idx / pt_idx [O] int32; observations grouped by point: ids
 *   pt_obs[pt_off[p] .. pt_off[p+1]).  Stops after num_iterations trials or num_success_iterations accepted steps.  state_out
 *   (device, 8 doubles, may be NULL): lambda, cost, candidate cost, accepted steps, done, last accepted, iterations, initial cost. */
int op_snavely_residual(const double *points, const double *cams, const double *features, const int64_t *pt_idx, const int64_t *cam_idx,
                        long long n, double *residual, void *stream);
int op_project_points(const double *xyz, const double *K, const double *RT, long long n, const double *ref2d, double *xy, double *depth,
                      double *dist_sum, void *stream);
int op_triangulate_points(const double *P1, const double *P2, const double *x1, const double *x2, long long n, double *X, void *stream);
size_t op_ba_workspace_bytes(int C, int P, int O);
int op_ba_lm(double *cams, double *points, const double *features, const int32_t *cam_idx, const int32_t *pt_idx, const int32_t *pt_off,
             const int32_t *pt_obs, int C, int P, int O, int num_iterations, int num_success_iterations, double lambda0, void *workspace,
             size_t workspace_bytes, double *state_out, void *stream);

/* ---- (4) batched PnP-RANSAC: eval_utils.py:18-42 (cv2.solvePnPRansac, EPnP, 5 px, 10000 it) ---
 * Correspondences of frame f are rows [f*stride, f*stride + count[f]) of pts2d/pts3d (float64,
 * converted to float32 inside as cv2 does).  samples: [n_samples_sets, H, 5] int32 minimal-sample
 * indices (host-generated cv::RNG stream); frame f uses sample set sample_set[f].
 * samples == sample_set == NULL: cv2's own stream is generated on the device, lazily per round.
 * Stage 1 solves one 5-point EPnP per hypothesis and counts inliers (float32 squared reprojection
 * error <= thr^2); stage 2 replays cv2's sequential accept / adaptive-niters rule, recomputes the
 * winner's inlier mask, refits EPnP on all inliers and converts to [R|t/scale].
 * Hypotheses are evaluated in rounds [0,32), [32,256), [256,1024), [1024,4096), ...; a frame stops taking part once
 * cv2's adaptive iteration bound, replayed over the rounds done, is reached (hyp_counts stay -1 beyond: same result as
 * evaluating all H, cv2 never looks at them).
 * Outputs: pose [F,3,4] float64, inlier_mask [F,stride] uint8, n_inliers [F] int32,
 * iters_run [F] int32, hyp_counts [F,H] int32 (diagnostic), hyp_pose [F,H,6] float64 (rvec,tvec). */
size_t op_pnp_workspace_bytes(int F, int stride, int H);
/* cv2's own minimal-sample stream, generated on the device: for frame f the first H index quintuples that
 * RANSACPointSetRegistrator::getSubset draws from cv::RNG((uint64)-1) for count[f] correspondences
 * (multiply-with-carry, redraw on duplicate).  samples [F,H,5] int32, sample_set [F] = 0..F-1. */
int op_pnp_cv_samples(const int32_t *count, int F, int H, int32_t *samples, int32_t *sample_set, void *stream);
int op_pnp_ransac(const double *pts2d, const double *pts3d, const int32_t *count, int F, int stride,
                  const double *K /* [F,3,3] */, const int32_t *samples, const int32_t *sample_set, int H,
                  double reproj_thr, double confidence, int max_iters, double scale,
                  double *pose, uint8_t *inlier_mask, int32_t *n_inliers, int32_t *iters_run,
                  int32_t *hyp_counts, double *hyp_pose, void *workspace, size_t workspace_bytes, void *stream);
/* Optional Levenberg-Marquardt polish of the poses op_pnp_ransac returned, on each frame's inliers (north_star item 4, "followed by a final
 * LM refinement").  NOT part of the parity path: the reference call (eval_utils.py:28-29) ends with cv2's EPnP on the inliers.  pts2d / pts3d /
 * count / K / scale as for op_pnp_ransac; inlier_mask [F,stride] and n_inliers [F] are its outputs; pose [F,3,4] is refined in place
 * (frames with fewer than 4 inliers are left alone).  rms (optional) [F,2]: reprojection RMS in pixels over the inliers before and after. */
int op_pnp_refine_lm(const double *pts2d, const double *pts3d, const int32_t *count, int F, int stride, const double *K,
                     const uint8_t *inlier_mask, const int32_t *n_inliers, double scale, int max_iters, double *pose, double *rms,
                     void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ONEPOSE_B200_H */
