/*
 * fcaf3d_b200 — C ABI of the B200-native FCAF3D hot path (libfcaf3d_b200.so).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*) and allocates nothing:
 *     outputs and scratch (`ws`, sized by the matching *_ws_bytes function) come from the caller;
 *   - return value: 0 = ok, <0 = error (-1 bad argument, -2 CUDA error, -3 workspace too small);
 *     b2_last_error() returns the message for the calling thread.  (The reference's ops instead
 *     terminate the process: mmdet3d/ops/pcdet_nms/src/iou3d_nms.cpp:14-38.)
 *   - coordinates are int32 rows (batch, x, y, z); rows of one map are grouped by ascending batch index;
 *     valid range: 0 <= batch < 1023, -131072 <= x,y,z < 131072 (packed 64-bit hash keys);
 *   - dtype codes: 0 = float32, 1 = bfloat16;  activation codes: 0 none, 1 ReLU, 2 ELU.
 *
 * "ME" below is MinkowskiEngine v0.5.4 (pinned at /root/reference docker/Dockerfile:24-32), whose C++/CUDA
 * backend these entry points replace for the calls the reference makes; ME is not vendored in the
 * reference tree, so the citations give the reference call sites.
 */
#ifndef FCAF3D_B200_H
#define FCAF3D_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ------------------------------------------------------------------------------- */
const char* b2_last_error(void);
int b2_version(void);
int b2_device_sm_count(void);
/* kernels launched by this library in this process so far (bench.py reports the count inside its timed region) */
long long b2_launch_count(void);

/* ---- coordinate manager (ME CoordinateMapManager; call sites single_stage_sparse.py:34-37,
 *      me_resnet.py:19-24,55-78, fcaf3d_neck_with_head.py:60-66,101,110-126) -------------------- */

/* points [n, dim] fp32 (xyz + features) -> coords[row0.., 4] = (batch, floor(xyz / voxel_size)),
 * feats[row0.., dim-3] = rest / feat_div.  single_stage_sparse.py:34-36 + ME.utils.batch_sparse_collate */
int b2_quantize_points(const float* pts, long long n, int dim, int batch, float voxel_size, float feat_div,
                       long long row0, int* coords, float* feats, void* stream);

size_t b2_scan_ws_bytes(long long n);
/* exclusive scan of a 0/1 byte mask: pos[i] = #set before i, *total = #set (order-preserving compaction) */
int b2_scan_mask(const uint8_t* mask, long long n, int* pos, int* total, void* ws, size_t ws_bytes, void* stream);

int b2_table_clear(void* table, long long cap, void* stream);
size_t b2_map_build_ws_bytes(long long n);
/* ME insert_and_map / stride: hash-insert n rows (optionally floored to multiples of `quant`), dedup with the
 * LOWEST input row as representative, rows in first-occurrence order (ME CPU rule).  table: cap slots of 16 B,
 * cap = power of two >= 2n.  Outputs: out_coords[*n_out,4], unique_index[*n_out], inverse_map[n],
 * *err != 0 if a coordinate is out of the packed range.  inverse_map == NULL (strided maps inside a network: nobody reads
 * it): the build is two kernels instead of three and unique_index may be NULL as well. */
int b2_map_build(const int* coords, long long n, int quant, void* table, long long cap, int* out_coords,
                 int* unique_index, int* inverse_map, int* n_out, int* err, void* ws, size_t ws_bytes, void* stream);
/* hash table over already-unique rows (value = row) */
int b2_map_table(const int* coords, long long n, void* table, long long cap, int* err, void* stream);
int b2_map_lookup(const void* table, long long cap, const int* query, long long m, int* rows, void* stream);
/* ME stride_region for MinkowskiGenerativeConvolutionTranspose(k=2,s=2): row 8*i+k = coords[i] + {0,half}^3,
 * k = ix + 2*iy + 4*iz.  fcaf3d_neck_with_head.py:60-66 */
int b2_map_children(const int* coords, long long n, int half, int* out_coords, void* stream);
/* offsets[b] = first row of batch b (offsets[num_batches] = n); *err = 2 if rows are not batch-grouped */
int b2_batch_offsets(const int* coords, long long n, int num_batches, int* offsets, int* err, void* stream);
/* same with the row count in device memory (*n_dev <= max_n): count + offsets can be fetched with one host read */
int b2_batch_offsets_dev(const int* coords, const int* n_dev, long long max_n, int num_batches, int* offsets, int* err,
                         void* stream);
size_t b2_map_union_ws_bytes(long long nb);
/* ME union_map (SparseTensor.__add__ with different keys, fcaf3d_neck_with_head.py:101): out = a U b, rows of
 * scene s = [a rows of s][b-only rows of s]; dst_a[na], dst_b[nb] give the output row of every input row. */
int b2_map_union(const void* a_table, long long a_cap, const int* a_coords, long long na, const int* a_off,
                 const int* b_coords, long long nb_rows, const int* b_off, int* out_coords, int* dst_a, int* dst_b,
                 int* n_out, void* ws, size_t ws_bytes, void* stream);
/* row compaction helpers (ME MinkowskiPruning, fcaf3d_neck_with_head.py:125) */
int b2_gather_coords(const int* coords, const uint8_t* mask, const int* pos, long long n, int* out, void* stream);
int b2_gather_rows(const void* in, const uint8_t* mask, const int* pos, long long n, int row_bytes, void* out,
                   void* stream);
int b2_index_rows(const void* in, const int* idx, long long n, int row_bytes, void* out, void* stream);
/* out[dst[i]] = in[i] (accumulate: +=); rows with dst[i] < 0 are skipped */
int b2_scatter_rows(const void* in, const int* dst, long long n, int c, int dtype, int accumulate, void* out,
                    void* stream);

/* ---- kernel maps + sparse convolution (ME kernel_map / ConvolutionForwardKernelGPU; call sites
 *      me_resnet.py:19-21,55-78, fcaf3d_neck_with_head.py:52,60-69,83-85) ------------------------ */

/* nbr[k*ld + o] = input row at coord_o + offset_k (or -1); offsets x-fastest (ME order), odd kernels centred;
 * tile_mask[o/128] bit k = some row of the 128-row tile has neighbour k.  ld = n_out rounded up to 128. */
int b2_kernel_map(const void* in_table, long long cap, const int* out_coords, long long n_out, int ksize, int step,
                  int* nbr, long long ld, unsigned int* tile_mask, void* stream);
/* The same table for a stride-2 convolution / pooling (out map = in coordinates floored to 2*step, kernel 2 or 3, offsets
 * step apart), built from the input rows by probing the OUT map's hash table (~3.4 probes per input row for kernel 3,
 * one for kernel 2, instead of kvol probes per output row); identical table and tile masks.  ld = n_out rounded to 128. */
int b2_kernel_map_strided(const void* out_table, long long cap, const int* in_coords, long long n_in, int ksize, int step,
                          int* nbr, long long ld, unsigned int* tile_mask, void* stream);
/* 3^3 stride-1 table of a children map (rows 8*i + c as made by b2_map_children, all children present) derived from
 * the parent map's 3^3 stride-1 table by index arithmetic, no hash probes; tile_mask is set to "all offsets" */
int b2_kernel_map_children(const int* parent_nbr, long long parent_ld, long long n_parent, int* nbr, long long ld,
                           unsigned int* tile_mask, void* stream);
/* Prune directly behind a children map (fcaf3d_neck_with_head.py:99-102, eval mode): the up-block's 3^3 convolution is
 * evaluated at the rows that survive the prune only.  b2_prune_plan: for the union's row maps dst_b (children row ->
 * union row, n_gen rows) and dst_a (backbone row -> union row, n_a rows), the prune mask and its exclusive scan pos:
 * gen_to_kept / a_to_kept = row in the pruned map or -1, kept_src[pruned row] = children row (every pruned row must be
 * a children row, i.e. union size == n_gen).  b2_kernel_map_children_pruned: from the parent's 3^3 table, nbr_up =
 * table from pruned rows into ALL children rows (the convolution's map), nbr_out = 3^3 table of the pruned map itself
 * (what b2_kernel_map gives on its coordinates), with exact tile masks; ld = n_keep rounded up to 128. */
int b2_prune_plan(const int* dst_b, long long n_gen, const int* dst_a, long long n_a, const uint8_t* mask, const int* pos,
                  int* gen_to_kept, int* a_to_kept, int* kept_src, void* stream);
int b2_kernel_map_children_pruned(const int* parent_nbr, long long parent_ld, const int* kept_src, const int* gen_to_kept,
                                  long long n_keep, long long ld, int* nbr_up, unsigned int* mask_up, int* nbr_out,
                                  unsigned int* mask_out, void* stream);
/* CUDA-core implicit GEMM (fp32 accumulate):
 *   out[o,:] = act((sum_k in[nbr[k][o],:] @ w[k] + bias) * scale + shift + residual[o,:]);  w: [kvol, cin, cout] fp32.
 * nbr == NULL: identity map (kvol == 1).  ch_mod > 0: epilogue channel = column % ch_mod. */
int b2_conv_cc(const void* in, int in_dtype, const float* w, const int* nbr, long long ld,
               const unsigned int* tile_mask, long long n_out, int cin, int cout, int kvol, const float* bias,
               const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
               int out_dtype, void* stream);
/* tcgen05 implicit GEMM, bf16 operands / fp32 TMEM accumulators; wt: [kvol, cout, cin] bf16 (K-major).
 * n_in = rows of the input feature matrix.  cin % 64 == 0, cout % 32 == 0.  Same epilogue as b2_conv_cc; out_dtype 0 or 1.
 * out_rows (may be NULL): position -> output row of a sparsity-sorted table from b2_kmap_sort (nbr / tile_mask are then the
 * sorted ones); results are bit-identical to the unsorted walk. */
int b2_conv_tc(const void* in_bf16, long long n_in, const void* wt_bf16, const int* nbr, long long ld,
               const unsigned int* tile_mask, const int* out_rows,
               long long n_out, int cin, int cout, int kvol, const float* bias, const float* scale, const float* shift,
               const void* residual, int act, int ch_mod, void* out, int out_dtype, void* stream);
/* b2_conv_tc / b2_head_fused run layers with many row tiles (>= 4 x SMs) on a persistent variant of the kernel
 * (csrc/conv_tcp.cu: CTAs stay resident and walk tiles, results bit-identical).  mode 0 = by size (default), 1 = never,
 * 2 = whenever the layer shape allows it; returns the previous mode (any other value only queries). */
int b2_tc_persist_mode(int mode);
/* The same kernel on fp32 rows and fp32 weights wt_f32 [kvol, cout, cin], read by the tensor cores as TF32 operands
 * (products of 10-bit mantissas, fp32 accumulation), fp32 output: the fp32-storage mode of the tensor-core path (ME's own
 * arithmetic is fp32 cuBLAS SGEMM; tolerance against it 1e-3 of the output's max, stated in tests/test_conv_gpu.py).
 * cin % 32 == 0, cout % 32 == 0. */
int b2_conv_tc_tf32(const void* in_f32, long long n_in, const void* wt_f32, const int* nbr, long long ld,
                    const unsigned int* tile_mask, const int* out_rows, long long n_out, int cin, int cout, int kvol,
                    const float* bias, const float* scale, const float* shift, const void* residual, int act, int ch_mod,
                    void* out, void* stream);

/* Sparsity-aware grouping of the rows of a neighbour table (no counterpart in ME 0.5.4, which runs one GEMM per kernel
 * offset; same idea as TorchSparse++'s sorted implicit GEMM): rows are ordered by their set of present offsets so that a
 * 128-row tile of b2_conv_tc only walks the offsets its rows use.  perm int32 [ld] (-1 in the padding),
 * nbr_sorted [kvol, ld], tile_mask_sorted [ld / 128]. */
size_t b2_kmap_sort_ws_bytes(long long n_out, int kvol);
int b2_kmap_sort(const int* nbr, long long ld, long long n_out, int kvol, int* perm, int* nbr_sorted,
                 unsigned int* tile_mask_sorted, void* ws, size_t ws_bytes, void* stream);

/* ---- pooling / normalisation / activation (ME MinkowskiMaxPooling, InstanceNorm, BatchNorm, ReLU, ELU;
 *      me_resnet.py:22-24,63; fcaf3d_neck_with_head.py:53-54,67-71) ------------------------------ */
int b2_maxpool(const void* in, int dtype, const int* nbr, long long ld, int kvol, long long n_out, int c, void* out,
               int* argmax, void* stream);
/* b2_affine_act (per-scene scale / shift [B, c], scene = out_coords[o][0], or scene 0 when NULL) followed by b2_maxpool
 * in one pass over bf16 rows (c % 8 == 0): the stem's InstanceNorm + ReLU + pooling, me_resnet.py:22-24 */
int b2_maxpool_affine(const void* in, const int* nbr, long long ld, int kvol, long long n_out, int c, const int* out_coords,
                      const float* scale, const float* shift, int act, void* out, void* stream);
size_t b2_segment_stats_ws_bytes(int nseg, int c);
/* per-segment (contiguous row ranges) channel mean and biased variance, two-pass, deterministic */
int b2_segment_stats(const void* x, int dtype, const int* seg_off, int nseg, int c, float* mean, float* var, void* ws,
                     size_t ws_bytes, void* stream);
/* MinkowskiInstanceNorm (me_resnet.py:22) folded to per-scene (scale, shift) [nseg, c] in one pass over the rows:
 * scale = weight / sqrt(var + eps), shift = bias - mean * scale with the per-segment biased variance (shifted sums,
 * deterministic).  dtype 0: c % 4 == 0, dtype 1: c % 8 == 0, and the 16-byte vectors of a row must divide 256. */
size_t b2_instnorm_ws_bytes(int nseg, int c);
int b2_instnorm_scale_shift(const void* x, int dtype, const int* seg_off, int nseg, int c, const float* weight,
                            const float* bias, float eps, float* scale, float* shift, void* ws, size_t ws_bytes,
                            void* stream);
/* y = act(x * scale[s,:] + shift[s,:] + residual), s = batch index of the row (coords) or 0 when coords == NULL */
int b2_affine_act(const void* x, int dtype, const int* coords, long long n, int c, const float* scale,
                  const float* shift, const void* residual, int act, void* y, void* stream);
int b2_cast(const void* x, int src_dtype, long long n, void* y, void* stream);

/* ---- head tail (fcaf3d_neck_with_head.py:110-126 _prune, :227-253 _get_bboxes_single, :256-279
 *      forward_single, :281-330 _bbox_pred_to_bbox, :332-374 _nms) ------------------------------- */
/* ME SparseTensor.features_at_coordinates: trilinear, absent corners contribute 0; query [m,4] fp32 (b,x,y,z) */
int b2_interp(const void* table, long long cap, const float* feats, int c, int tensor_stride, const float* query,
              long long m, float* out, void* stream);
/* mask[i] = 1 for the k largest scores of each segment (ties at the k-th value: lowest rows first) */
int b2_topk_mask(const float* scores, const int* seg_off, int nseg, int k, uint8_t* mask, void* stream);
/* same, with the total row count known to the host: segments of >= 32768 rows on average are selected by a cluster of 8
 * CTAs per segment (histograms merged through distributed shared memory); masks are identical to b2_topk_mask */
int b2_topk_mask_n(const float* scores, const int* seg_off, int nseg, int k, long long n_total, uint8_t* mask,
                   void* stream);
/* feat [n, feat_ld] (first 1+n_reg+n_cls columns used) -> centerness[n], bbox_pred[n,n_reg] (first 6: exp(*scale * x), scale = the level's
 * learnable Scale on the device), cls[n,n_cls], prune_score[n] = max class logit,
 * max_score[n] = max_c sigmoid(cls_c)*sigmoid(centerness) (optional) */
int b2_head_post(const float* feat, long long n, int feat_ld, int n_reg, int n_cls, const float* scale,
                 float* centerness, float* bbox_pred, float* cls, float* prune_score, float* max_score, void* stream);
/* forward_single (fcaf3d_neck_with_head.py:256-279) in ONE launch for bf16 rows: the three 1x1 head convolutions as one tcgen05
 * GEMM [n, cin] x wt_bf16 [cols = 32 (padded), cin] + bias [32] whose epilogue writes what b2_head_post writes (bit-identical):
 * no [n, 32] logit tensor goes to memory. */
int b2_head_fused(const void* in_bf16, long long n, const void* wt_bf16, int cin, int cols, const float* bias, int n_reg,
                  int n_cls, const float* level_scale, float* centerness, float* bbox_pred, float* cls, float* prune_score,
                  float* max_score, void* stream);
/* yaw_mode: 0 axis-aligned, 1 naive, 2 sin-cos, 3 fcaf3d (Mobius); out [n,7] */
int b2_decode_boxes(const float* points, const float* bbox_pred, long long n, int n_reg, int yaw_mode, float* out,
                    void* stream);
int b2_emit_candidates(const int* coords, const float* centerness, const float* bbox_pred, const float* cls,
                       const uint8_t* mask, const int* pos, const int* seg_off, const int* dst_base, long long n,
                       int n_reg, int n_cls, int yaw_mode, float voxel_size, float* cand_boxes, float* cand_scores,
                       void* stream);
int b2_class_segments(const float* cand_boxes, const float* cand_scores, const int* cand_off, int nscene, int n_cls,
                      long long total, float score_thr, uint8_t* flags, int* pos, int* n_flagged, float* seg_boxes,
                      float* seg_scores, int* seg_src, int* seg_off, void* ws, size_t ws_bytes, void* stream);
int b2_collect_dets(const float* sorted_boxes, const float* seg_scores, const int* order, const int* seg_off, int nseg,
                    int n_cls, const uint8_t* keep, const int* pos, long long n, float* out_boxes, float* out_scores,
                    int* out_labels, int* out_scene, void* stream);

/* ---- BEV NMS (mmdet3d/ops/pcdet_nms: iou3d_nms.h:9-12 nms_gpu / nms_normal_gpu / boxes_iou_bev_gpu /
 *      boxes_overlap_bev_gpu, kernels iou3d_nms_kernel.cu:104-372, host reduce iou3d_nms.cpp:116-132) ---- */
/* HOST helper: tile_off_host[nseg+1], mask_off_host[nseg+1] (8-byte words), *mask_words, *max_seg */
int b2_nms_plan(const int* seg_off_host, int nseg, int* tile_off_host, long long* mask_off_host, long long* mask_words,
                int* max_seg);
/* all segments in one launch per stage; keep[n] in (per-segment) sorted order, counts[nseg].
 * sort == 0 reproduces nms_gpu / nms_normal_gpu on pre-sorted boxes (sorted_boxes/order unused). */
int b2_nms_batched(const float* boxes, const float* scores, const int* seg_off, const int* tile_off,
                   const long long* mask_off, int nseg, int total_tiles, int max_seg, float thresh, int rotated,
                   int sort, float* sorted_boxes, int* order, unsigned long long* mask, uint8_t* keep, int* counts,
                   void* stream);
/* same result as b2_nms_batched, computed block by block so that only rows of KEPT boxes are ever evaluated (the
 * reference's host loop reads no others, iou3d_nms.cpp:121-131) and no mask is written to global memory */
size_t b2_nms_greedy_ws_bytes(long long n, int rotated);
int b2_nms_greedy_batched(const float* boxes, const float* scores, const int* seg_off, int nseg, long long n,
                          int max_seg, float thresh, int rotated, int sort, float* sorted_boxes, int* order, void* ws,
                          size_t ws_bytes, uint8_t* keep, int* counts, void* stream);
/* mode 0 = boxes_overlap_bev_gpu, 1 = boxes_iou_bev_gpu, 2 = axis-aligned BEV IoU (iou_normal) */
int b2_boxes_pair_matrix(const float* a, long long na, const float* b, long long nb, int mode, float* out,
                         void* stream);
int b2_keep_list(const uint8_t* keep, const int* pos, long long n, long long* out, void* stream);
/* (No counterpart in the reference.)  Greedy BEV NMS of one pre-sorted box list that allocates nothing at all:
 * for callers that may not allocate device memory at call time (a frame loop that fixed its GPU budget at start-up,
 * a host runtime that owns the GPU heap).  The mask scratch that b2_nms_batched takes from the caller lives in a
 * ~0.5 MB array inside the library image (held on every GPU that loads the library until exit), hence the limit
 * n <= b2_nms_single_noalloc_max_boxes() (2048); larger n returns -1.  boxes_sorted [n,7], keep [n] (1 = kept),
 * count [1]: device memory from the caller.  The first call on a device resolves the array's device address — make
 * it at start-up, before any stream capture.  Calls on one device share the array: overlapping calls race. */
int b2_nms_single_noalloc_max_boxes(void);
int b2_nms_single_noalloc(const float* boxes_sorted, long long n, float thresh, int rotated, uint8_t* keep, int* count,
                          void* stream);

/* ---- device-side data path (reference: numpy pipeline datasets/pipelines/transforms_3d.py) ----------- */
/* IndoorPointSample (:838-880) + RandomFlip3D (:144-160) + GlobalRotScaleTrans (:541-648) in one pass per scene:
 * out[i] = T(in[choices[i]]), T = flip signs, rotation about z by `angle` (xyz @ rot_mat_T as base_points.py:139-179),
 * uniform scale, translation; attribute columns (colour) are copied.  choices int64 [n_out] device (NULL = identity).
 * err int32 [1] device flag, zeroed by the caller: 1 = index out of range.  Random numbers are drawn by the caller. */
int b2_augment_points(const float* in, long long n_in, int dim, const long long* choices, long long n_out, int flip_h,
                      int flip_v, float angle, float scale, float tx, float ty, float tz, float* out, int* err,
                      void* stream);
/* the same maps on ground-truth boxes (cx, cy, cz, dx, dy, dz, yaw), gravity centre (depth_box3d.py:119-208,
 * base_box3d.py:149-222): yaw -> pi - yaw / -yaw under flips, yaw -= angle; without yaw the x / y sizes become the extents
 * of the rotated footprint */
int b2_augment_boxes(const float* boxes, int n, int with_yaw, int flip_h, int flip_v, float angle, float scale, float tx,
                     float ty, float tz, float* out, void* stream);

/* ---- training path ----------------------------------------------------------------------------- */
/* ME ConvolutionBackwardKernelGPU, weight part: dw[k,ci,co] += sum_o in[nbr[k][o],ci] * dout[o,co] (fp32 atomics;
 * dw zero-initialised by the caller).  The input gradient is b2_conv_cc / b2_conv_tc over the transposed map with
 * transposed weights (host side: fcaf3d_b200/autograd.py). */
int b2_conv_wgrad(const void* in, int a_dtype, const void* dout, int g_dtype, const int* nbr, long long ld,
                  long long n_out, int cin, int cout, int kvol, float* dw, void* stream);
/* ME max-pool backward: dx[argmax[o,c], c] += dy[o,c]; dx fp32, zero-initialised by the caller */
/* bf16 operand layouts of a convolution kernel w [kvol, cin, cout] fp32 for b2_conv_tc, one pass: wt_fwd [kvol, cout, cin]
 * (forward) and wt_bwd [kvol, cin, cout] with the offsets reversed (input gradient over the transposed map); either may
 * be NULL.  Replaces the per-step torch transposes of ME's kernel tensor (MinkowskiConvolution.kernel, App. A.4). */
int b2_pack_conv_weights(const float* w, int kvol, int cin, int cout, void* wt_fwd, void* wt_bwd, void* stream);
int b2_maxpool_bwd(const void* dy, int dtype, const int* argmax, long long n_out, int c, float* dx, void* stream);
/* Training-mode MinkowskiBatchNorm (= nn.BatchNorm1d on the feature rows, ME App. A.6) fused with activation / residual:
 * statistics of x [n, c] (dtype 0 f32 / 1 bf16, c % 8 == 0) -> mean, invstd, folded scale / shift (apply with
 * b2_affine_act); running statistics and num_batches_tracked updated like torch (momentum, unbiased variance). */
size_t b2_bn_ws_bytes(int c);
int b2_bn_train_stats(const void* x, int dtype, long long n, int c, const float* weight, const float* bias,
                      float* running_mean, float* running_var, long long* num_batches_tracked, float momentum, float eps,
                      float* mean, float* invstd, float* scale, float* shift, void* ws, size_t ws_bytes, void* stream);
/* backward of y = act(bn(x) [+ residual]) (act: 0 none, 1 ReLU, 2 ELU; derivative taken from y): sums f32 [2, c] =
 * (d bias, d weight), dx [n, c], dres [n, c] (gradient of the residual branch) or NULL */
int b2_bn_act_bwd(const void* x, const void* dy, const void* y, int dtype, long long n, int c, int act, const float* mean,
                  const float* invstd, const float* weight, float* sums, void* dx, void* dres, void* ws, size_t ws_bytes,
                  void* stream);
/* Fcaf3DAssigner.assign for one scene (fcaf3d_neck_with_head.py:387-466): points [P,3] of all levels (level_off[L+1]),
 * boxes [B,7] = (gravity centre, size, yaw), labels_in int64 [B] -> centerness_t [P], box_t [P,7], labels_out int64 [P]
 * (-1 = background).  Scratch: counts int32 [L*B], best_level int32 [B], thr f32 [B]. */
int b2_fcaf3d_assign(const float* points, long long P, const int* level_off, int L, const float* boxes, int B,
                     const long long* labels_in, int limit, int topk, int* counts, int* best_level, float* thr,
                     float* centerness_t, float* box_t, long long* labels_out, void* stream);
/* The same assignment for ALL scenes of a batch in three launches (removes the reference's python loop over
 * scenes, fcaf3d_neck_with_head.py:128-141).  points [P,3] level-major (rows of each level's coordinate map, scenes grouped
 * inside a level); seg int32 [L*S][2] = row range of (level, scene) at index level*S+scene; boxes [Btot,7] of all scenes
 * with box_off int32 [S+1]; extra outputs scene_of int32 [P] and n_pos int32 [S] (positives per scene). */
int b2_fcaf3d_assign_batched(const float* points, long long P, const int* seg, int S, int L, const float* boxes,
                             const int* box_off, int Btot, const long long* labels_in, int limit, int topk, int* counts,
                             int* best_level, float* thr, float* centerness_t, float* box_t, long long* labels_out,
                             int* scene_of, int* n_pos, void* stream);
/* mmcv.ops.sigmoid_focal_loss element-wise forward and d/dlogit (label outside [0,c) = all-negative row) */
int b2_sigmoid_focal(const float* logits, const long long* labels, long long n, int c, float gamma, float alpha,
                     float* loss, float* grad, void* stream);
/* aligned-pair axis-aligned 3-D IoU (iou3d_calculator.py:279-312 through iou3d_loss.py:21-35) and d iou / d pred */
int b2_aa_iou3d(const float* pred, const float* tgt, long long n, int stride, float* iou, float* dpred, void* stream);
/* Rotated_IoU cuda_op.sort_v (box_intersection_2d.py:147): vertices [n,24,2], mask [n,24] u8, num_valid [n] -> idx [n,9] */
int b2_sort_vertices(const float* vertices, const uint8_t* mask, const int* num_valid, long long n, int* idx,
                     void* stream);
/* Rotated 3-D IoU of aligned pairs of (x, y, z, w, l, h, alpha) boxes and d iou / d pred [n,7] (NULL = value only), one
   launch: cal_iou_3d (oriented_iou_loss.py:86-109 -> :38-58, :6-35) -> oriented_box_intersection_2d
   (box_intersection_2d.py:13-183) with the angular sort of the un-vendored cuda_op.sort_v done in-thread.  The gradient
   is the one torch autograd gives for that graph (iou3d_loss.py:11-13 wraps it as 1 - iou). */
int b2_rot_iou3d(const float* pred, const float* tgt, long long n, float* iou, float* dpred, void* stream);
/* The head's regression loss in one launch: decode bbox_pred [n, n_reg] at points [n,3] (_bbox_pred_to_bbox,
   fcaf3d_neck_with_head.py:281-330; yaw_mode 0 none, 1 naive, 2 sin-cos, 3 fcaf3d), IoU with tgt [n,7] (rotated != 0:
   as b2_rot_iou3d; 0: axis-aligned, iou3d_loss.py:21-35 -> iou3d_calculator.py:279-312) -> iou [n] and
   d iou / d bbox_pred [n, n_reg]. */
int b2_box_iou_decoded(const float* points, const float* bbox_pred, int n_reg, int yaw_mode, const float* tgt,
                       int rotated, long long n, float* iou, float* dbbox_pred, void* stream);
/* element-wise binary cross entropy with logits and d/dx (fcaf3d_neck_with_head.py:191-193 -> mmdet 2.14
   CrossEntropyLoss(use_sigmoid=True) -> F.binary_cross_entropy_with_logits) */
int b2_bce_logits(const float* x, const float* t, long long n, float* loss, float* grad, void* stream);
/* Optimiser of the reference recipe (configs/fcaf3d/fcaf3d.py:30-31: AdamW + grad_clip max_norm 10; mmcv OptimizerHook ->
 * torch clip_grad_norm_ + torch.optim.AdamW) over ALL parameter tensors at once.  table: n records of 5 x int64 (param,
 * grad, exp_avg, exp_avg_sq device pointers, element count); chunks: n_chunks int32 pairs (tensor index, chunk index) with
 * b2_opt_chunk_elems() elements per chunk.  b2_grad_sqnorm: deterministic squared L2 norm of every gradient element
 * (partial: n_chunks floats of scratch).  b2_adamw_step: gradient scaled by min(1, max_norm / (sqrt(*sqnorm) + 1e-6))
 * (max_norm <= 0 or sqnorm NULL: unscaled), decoupled weight decay, moments, bias-corrected update of 1-based `step`. */
int b2_opt_chunk_elems(void);
int b2_grad_sqnorm(const void* table, const int* chunks, int n_chunks, float* partial, float* sqnorm_out, void* stream);
int b2_adamw_step(const void* table, const int* chunks, int n_chunks, double lr, double beta1, double beta2, double eps,
                  double wd, long long step, double max_norm, const float* sqnorm, void* stream);

/* ---- native executor of one fused inference step ------------------------------------------------------------------
 * What the reference does per batch through python - SingleStageSparse3DDetector.simple_test -> extract_feat ->
 * MEResNet3D.forward -> Fcaf3DNeckWithHead.forward / get_bboxes (single_stage_sparse.py:32-59, me_resnet.py:66-99,
 * fcaf3d_neck_with_head.py:94-126,205-253,256-279,332-374), eval mode - as ONE host call that enqueues the entry points
 * above in C++ (bit-identical results to calling them one by one; the python host needed as long per step as the GPU).
 * b2_engine_create (set-up, not per step): parses the weight table `desc` (int64 words, layout documented at the
 * definition in csrc/engine.cu and written by fcaf3d_b200/engine.py; pointers are device pointers owned by the caller, which
 * must outlive the engine) and fl = {voxel_size, score_thr, iou_thr, InstanceNorm eps}; allocates a small pinned staging
 * ring and events.  Returns NULL on error.  One engine per device; calls on one engine must not overlap on the host.
 * b2_engine_infer (per step, allocates nothing): points_host / n_points_host are HOST arrays of n_scenes device pointers
 * ([n_i, dim] fp32 rows xyz + colour) and row counts; dtype = activation storage (0 fp32, 1 bf16, 2 fp32 with the wide
 * convolutions as TF32 tensor-core operands: the table's `wt` pointers then hold fp32 [kvol, cout, cin]); every intermediate is
 * taken from `arena` (-3 if too small; *arena_used_host reports the bytes used); detections come back in the caller's
 * buffers of out_cap >= b2_engine_out_rows() rows: out_boxes [.,7], out_scores, out_labels, out_scene, grouped by
 * (scene, class), *out_count (device) of them valid; *n_rows_host = rows the tail worked on.  The call blocks on the few
 * row-count reads of the coordinate maps (each queued behind GPU work that was enqueued earlier).
 * tail_stream (NULL or == stream: everything on `stream`): the top-k / decode / NMS tail - small launches on a handful
 * of SMs - is enqueued on that second stream behind an event, so the NEXT call's first kernels on `stream` overlap it; the
 * outputs are then complete when tail_stream reaches the point the call left it at, and the caller must not hand the same
 * arena to a later call before that (fcaf3d_b200/engine.py alternates two arenas and orders them with events).
 * Coordinate stream: the engine owns a third stream for everything that depends on coordinates only (hash builds, kernel
 * maps); the feature kernels on `stream` wait for it by events, never the other way round, so the coordinate work of the
 * next call runs while the GPU is still busy with this call's convolutions.  That stream must know when it may start:
 * inputs_ready_event (a cudaEvent_t, e.g. the end of the host->device copy) or inputs_ready != 0 (the points are complete
 * already) - with neither it waits for everything queued on `stream` so far; arena_free_event (cudaEvent_t or NULL): the
 * event after which `arena` may be overwritten.  B2_ENGINE_SPLIT=0 keeps all of it on `stream`. */
void* b2_engine_create(const long long* desc, int n_words, const double* fl);
int b2_engine_destroy(void* engine);
long long b2_engine_out_rows(void* engine, int n_scenes);
int b2_engine_infer(void* engine, const float* const* points_host, const long long* n_points_host, int n_scenes, int dtype,
                    void* arena, size_t arena_bytes, float* out_boxes, float* out_scores, int* out_labels, int* out_scene,
                    int* out_count, long long out_cap, long long* n_rows_host, size_t* arena_used_host, void* stream,
                    void* tail_stream, void* inputs_ready_event, int inputs_ready, void* arena_free_event);

#ifdef __cplusplus
}
#endif
#endif /* FCAF3D_B200_H */
