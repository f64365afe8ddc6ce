/*
 * jtsm_b200.h -- C ABI of libjtsm_b200.so: the B200 (sm_100a) implementation of
 * JTSM's per-proposal hot path (RoI pooling fwd/bwd, WSDDN MIL head, OICR
 * refinement label assignment + weighted losses, NMS).
 *
 * Conventions (every entry point):
 *   - plain pointers + sizes only; all pointers are DEVICE pointers unless the
 *     parameter name ends in `_host`;
 *   - `stream` is a cudaStream_t passed as void*; launches are stream-ordered,
 *     never synchronise, never allocate; scratch comes from a caller-provided
 *     workspace whose size is returned by the matching *_workspace_bytes();
 *   - return value: JTSM_OK (0) or a JTSM_ERR_* code; launch failures return
 *     JTSM_ERR_CUDA_BASE + cudaError_t.  Nothing throws.
 *   - `dtype` is the element type of feature / gradient tensors
 *     (JTSM_F32 / JTSM_BF16 / JTSM_F16); boxes, scores, losses are fp32;
 *     labels / indices are int64 where the reference exposes int64
 *     (gt_classes, keep indices) and int32 for argmax (reference: at::kInt).
 *   - RoI dtype rule: `rois` is ALWAYS fp32 at this boundary, whatever `dtype` the features have, and the
 *     kernels use the coordinate values as passed.  The reference's ROIAlign module rounds the coordinates
 *     through the feature dtype before its op sees them (`rois.to(dtype=input.dtype)`,
 *     detectron2/layers/roi_align.py:108); a binding that wants that behaviour does the same cast on its
 *     side and widens the result back to fp32 (exact) -- jtsm_b200/layers/roi_align.py does.
 *   - class-column limits: the MIL / refinement kernels hold a row's class columns in registers and
 *     support num_classes <= 256 (JTSM_ERR_INVALID_ARGUMENT above that).
 *
 * Each function cites the reference interface it replaces; paths are relative
 * to the reference checkout (shenyunhang/JTSM).
 */
#ifndef JTSM_B200_H_
#define JTSM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define JTSM_API __declspec(dllexport)
#else
#define JTSM_API __attribute__((visibility("default")))
#endif

#define JTSM_ABI_VERSION 1

enum {
  JTSM_OK = 0,
  JTSM_ERR_INVALID_ARGUMENT = 1,
  JTSM_ERR_UNSUPPORTED = 2,
  JTSM_ERR_WORKSPACE_TOO_SMALL = 3,
  JTSM_ERR_STATIC_LIMIT = 4, /* size beyond the fixed maximum of an allocation-free (*_static) entry point */
  JTSM_ERR_CUDA_BASE = 1000 /* + cudaError_t */
};

enum { JTSM_F32 = 0, JTSM_BF16 = 1, JTSM_F16 = 2 };
/* storage type of the saved argmax tensor */
enum { JTSM_ARGMAX_I32 = 0, JTSM_ARGMAX_U16 = 1 };

JTSM_API int jtsm_abi_version(void);
JTSM_API const char* jtsm_error_string(int code);
/* Number of kernels this library has launched in this process (all streams).
 * bench.py reports the delta over the timed region as `gpu_launches`. */
JTSM_API unsigned long long jtsm_launch_count(void);
/* Tuning knobs for A/B measurements (kernel variant selection). Unknown keys
 * return JTSM_ERR_INVALID_ARGUMENT. Not part of the reference surface. */
JTSM_API int jtsm_set_tuning(const char* key, int value);
JTSM_API int jtsm_get_tuning(const char* key, int* value);

/* ------------------------------------------------------------------------
 * RoIPool (max pooling with integer bin edges).
 * Replaces torchvision.ops.roi_pool as constructed by
 *   projects/WSL/wsl/modeling/poolers.py:183-186 and
 *   detectron2/modeling/poolers.py:161-164
 * (same arithmetic as the in-tree copy
 *   projects/WSL/wsl/layers/csrc/ROILoopPool/ROILoopPool_cpu.cpp:26-79,83-123).
 *
 * input   [batch, channels, height, width]  NCHW contiguous, `dtype`
 * rois    [num_rois, 5] fp32 (batch_index, x1, y1, x2, y2)
 * output  [num_rois, channels, pooled_h, pooled_w] `dtype`
 * argmax  same shape; JTSM_ARGMAX_I32 (-1 = empty) or JTSM_ARGMAX_U16
 *         (0xFFFF = empty; requires height*width < 65535)
 * row_scale  optional [num_rois] fp32: output row r is multiplied by
 *         row_scale[r] (fuses `box_features * (objectness_logits + 1)`,
 *         roi_heads_oicr.py:393-394); backward applies the same factor.
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_roi_pool_workspace_bytes(int batch, int num_rois,
                                              int pooled_h, int pooled_w);
JTSM_API int jtsm_roi_pool_forward(
    const void* input, const float* rois, int dtype, int batch, int channels,
    int height, int width, int num_rois, float spatial_scale, int pooled_h,
    int pooled_w, const float* row_scale, void* output, void* argmax,
    int argmax_dtype, void* workspace, size_t workspace_bytes, void* stream);
JTSM_API size_t jtsm_roi_pool_backward_workspace_bytes(int dtype, int batch,
                                                       int channels, int height,
                                                       int width);
/* grad_output may be non-contiguous: element strides (in elements) are passed
 * explicitly, as the reference's backward does (ROILoopPool_cuda.cu:358-361).
 * grad_input [batch, channels, height, width] is fully overwritten. */
JTSM_API int jtsm_roi_pool_backward(
    const void* grad_output, long long stride_n, long long stride_c,
    long long stride_h, long long stride_w, const float* rois,
    const void* argmax, int argmax_dtype, int dtype, int batch, int channels,
    int height, int width, int num_rois, int pooled_h, int pooled_w,
    const float* row_scale, void* grad_input, void* workspace,
    size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------
 * RoIAlign (bilinear average pooling, adaptive sampling grid).
 * Replaces roi_align_forward / roi_align_backward of detectron2._C
 *   (detectron2/layers/csrc/ROIAlign/ROIAlign.h:54-61,86-97, bound at vision.cpp:96-97, called from
 *   detectron2/layers/roi_align.py:31-33,44-56) and torchvision.ops.roi_align (roi_align.py:5-15).
 * Same argument meaning: sampling_ratio <= 0 -> ceil(roi_size / pooled) samples per bin axis;
 * aligned != 0 -> half-pixel shift and no minimum RoI size.
 * workspace: jtsm_roi_align_workspace_bytes(dtype, batch, channels, height, width, num_rois, backward).
 * Backward takes explicit grad_output strides (ROIAlign_cpu.cpp:478-482) and overwrites grad_input.
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_roi_align_workspace_bytes(int dtype, int batch, int channels, int height,
                                               int width, int num_rois, int backward);
JTSM_API int jtsm_roi_align_forward(const void* input, const float* rois, int dtype, int batch,
                                    int channels, int height, int width, int num_rois,
                                    float spatial_scale, int pooled_h, int pooled_w,
                                    int sampling_ratio, int aligned, void* output, void* workspace,
                                    size_t workspace_bytes, void* stream);
JTSM_API int jtsm_roi_align_backward(const void* grad_output, long long stride_n,
                                     long long stride_c, long long stride_h, long long stride_w,
                                     const float* rois, int dtype, int batch, int channels,
                                     int height, int width, int num_rois, float spatial_scale,
                                     int pooled_h, int pooled_w, int sampling_ratio, int aligned,
                                     void* grad_input, void* workspace, size_t workspace_bytes,
                                     void* stream);
/* Deterministic backward (opt-in): a segmented reduction instead of atomics -- channel-independent tap list, stable
 * radix sort by map cell, one fixed-order sum per (channel, cell).  Run-to-run bit-identical and bit-equal to the
 * reference's serial CPU backward (ROIAlign_cpu.cpp:168-263: same operands (g * w / count), same order).
 * sampling_ratio > 0 only (JTSM_ERR_UNSUPPORTED otherwise: the adaptive grid has no a-priori tap count).
 * grad_input is fully overwritten. */
JTSM_API size_t jtsm_roi_align_backward_det_workspace_bytes(int batch, int channels, int height,
                                                            int width, int num_rois, int pooled_h,
                                                            int pooled_w, int sampling_ratio);
JTSM_API int jtsm_roi_align_backward_det(const void* grad_output, long long stride_n,
                                         long long stride_c, long long stride_h,
                                         long long stride_w, const float* rois, int dtype,
                                         int batch, int channels, int height, int width,
                                         int num_rois, float spatial_scale, int pooled_h,
                                         int pooled_w, int sampling_ratio, int aligned,
                                         void* grad_input, void* workspace,
                                         size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------
 * ROILoopPool: three RoIPool-style max-pools per RoI -- the RoI, the "frame" (RoI minus the inner
 * rectangle shrunk by context_ratio) and the "context" (outer rectangle grown by context_ratio,
 * minus the RoI).  Replaces wsl._C.roi_loop_pool_forward (projects/WSL/wsl/layers/csrc/
 * ROILoopPool/ROILoopPool.h:48-53, ROILoopPool_cuda.cu:10-204; context_ratio = 1.8 at :309).
 * output / argmax (int32, -1 = none) are [3 * num_rois, channels, pooled_h, pooled_w]: rows [0,R) RoI,
 * [R,2R) frame, [2R,3R) context.  maxval starts at 0 as in the reference (input assumed >= 0).
 * Backward (ROILoopPool.h:67-77, _cuda.cu:207-248) = jtsm_roi_pool_backward over 3R rows with
 * rois repeated three times.
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_roi_loop_pool_workspace_bytes(int batch, int height, int width,
                                                    int num_rois, int pooled_h, int pooled_w);
JTSM_API int jtsm_roi_loop_pool_forward(const void* input, const float* rois, int dtype, int batch,
                                        int channels, int height, int width, int num_rois,
                                        float spatial_scale, float context_ratio, int pooled_h,
                                        int pooled_w, void* output, int* argmax, void* workspace,
                                        size_t workspace_bytes, void* stream);
/* ------------------------------------------------------------------------
 * MOIPool: RoI max-pool restricted to cells covered by a superpixel that is switched on for the RoI.
 * Replaces wsl._C.moi_pool_forward (projects/WSL/wsl/layers/csrc/MOIPool/MOIPool.h:52-59,
 * MOIPool_cuda.cu:139-215 mask, :218-294 pool).  oh_labels int32 [num_rois, len_labels],
 * superpixels int32 [batch, height_sp, width_sp].  Returns output AND argmax (int32; consumed by
 * roi_heads_jtsm.py:615-630).  Backward (MOIPool.h:85-95) = jtsm_roi_pool_backward.
 * Fused epilogue of the JTSM heads (roi_heads_jtsm.py:615-633), optional: out = (out * mask_scale[r]) * row_scale[r]
 * with mask_scale[r] = P*P / (#bins of RoI r whose arg-max is not -1, + 1) when `mask_scale` != 0 and row_scale =
 * objectness_logits + 1 (NULL = 1).  scale_out float[3 * num_rois] (required when either is on): mask scales,
 * row scales, products -- pass scale_out + 2*num_rois as row_scale to jtsm_roi_pool_backward.
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_moi_pool_workspace_bytes(int batch, int height, int width, int num_rois,
                                              int pooled_h, int pooled_w);
JTSM_API int jtsm_moi_pool_forward(const void* input, const float* rois, const int* oh_labels,
                                   int len_labels, const int* superpixels, int height_sp,
                                   int width_sp, int dtype, int batch, int channels, int height,
                                   int width, int num_rois, float spatial_scale, int pooled_h,
                                   int pooled_w, const float* row_scale, int mask_scale,
                                   float* scale_out, void* output, int* argmax, void* workspace,
                                   size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------
 * WSDDN two-stream MIL head: per image  scores = softmax(C, dim=1) * softmax(D, dim=0),
 * image scores = clamp(sum_r scores, 1e-6, 1-1e-6), loss = BCE(image scores, one-hot).
 * Replaces the eager composition in
 *   projects/WSL/wsl/modeling/roi_heads/fast_rcnn_wsddn.py:571-584 (WSDDNOutputLayers.forward),
 *   :364-379 (predict_probs_img) and :346-362 (binary_cross_entropy_loss; mean_loss=0 -> sum / N).
 *
 * cls_logits, det_logits, scores  [num_rois, num_classes] fp32 (num_classes <= 256)
 * seg_offsets   [num_images + 1] int32, DEVICE: rows of image n are [seg[n], seg[n+1])
 * labels_onehot [num_images, num_classes] fp32 or NULL (no loss)
 * img_scores    [num_images, num_classes] clamped image scores (= pred_class_img_logits)
 * col_stats     [num_images, num_classes, 3] (col max of D, col sum-exp, unclamped image sum): saved for backward
 * loss          [1] or NULL;  grad_img [num_images, num_classes] d loss / d (unclamped image sum) or NULL
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_wsddn_workspace_bytes(int num_images);
JTSM_API int jtsm_wsddn_forward(const float* cls_logits, const float* det_logits,
                                const int* seg_offsets, int num_images, int num_rois,
                                int num_classes, const float* labels_onehot, int mean_loss,
                                float* scores, float* img_scores, float* col_stats, float* loss,
                                float* grad_img, void* workspace, size_t workspace_bytes,
                                void* stream);
/* d(cls_logits), d(det_logits) for upstream gradient
 *   g[r,k] = grad_scores[r,k] (NULL = 0) + grad_img[n(r),k] * (*grad_loss) (NULL pointers = 0 / 1). */
JTSM_API int jtsm_wsddn_backward(const float* cls_logits, const float* det_logits,
                                 const float* scores, const float* col_stats,
                                 const int* seg_offsets, int num_images, int num_rois,
                                 int num_classes, const float* grad_scores, const float* grad_img,
                                 const float* grad_loss, float* grad_cls, float* grad_det,
                                 void* stream);

/* ------------------------------------------------------------------------
 * OICR refinement, one branch.
 * jtsm_refine_assign = get_pgt_top_k (top_k = 1; roi_heads_oicr.py:675-811) +
 *   label_and_sample_proposals (roi_heads.py:263-370: pairwise_iou boxes.py:369-391, Matcher
 *   matcher.py:61-103, _sample_proposals roi_heads.py:222-254) + the weight preparation of
 *   OICROutputs.__init__ (fast_rcnn_oicr.py:243-247), per image, one launch.
 *
 * prev_scores  [num_rois, score_stride]: class k in column k (WSDDN scores: stride K; a previous
 *              branch's softmax: stride K+1)
 * prev_boxes   [num_rois, boxes_per_row, 4]: boxes_per_row = 1 (proposal boxes / class-agnostic)
 *              or num_classes (class-specific predict_boxes output)
 * proposals    [num_rois, 4]
 * img_cls_offsets [num_images+1] int32 DEVICE, img_cls_ids int64 DEVICE: gt_classes_img_int, ragged
 * img_scores   [num_images, num_classes]: pred_class_img_logits (pseudo-gt weights)
 * Matcher: thresholds iou_t0 (< iou_t1 if num_thresholds == 2), labels label0..label2 in {-1,0,1}
 * outputs: pgt_* [total seeds] (seed g of image n at img_cls_offsets[n] + g);
 *   matched_idx int64 [R] (seed index within the image), matched_iou [R] or NULL,
 *   gt_classes int64 [R] (class | num_classes = background | -1 = ignore), gt_weights, gt_scores [R],
 *   gt_boxes [R,4], loss_weights [R] (gt_weights with ignored rows zeroed),
 *   counters int32 [num_images*3 + 1]: (fg, bg, ignore) per image, then #(loss_weights > 1e-12).
 * ------------------------------------------------------------------------ */
JTSM_API int jtsm_refine_assign(
    const float* prev_scores, int score_stride, const float* prev_boxes, int boxes_per_row,
    const float* proposals, const int* seg_offsets, int num_images, int num_rois, int num_classes,
    const int* img_cls_offsets, const long long* img_cls_ids, int max_classes_per_image,
    const float* img_scores, float iou_t0, float iou_t1, int label0, int label1, int label2,
    int num_thresholds, float* pgt_boxes, long long* pgt_classes, float* pgt_scores,
    float* pgt_weights, long long* matched_idx, float* matched_iou, long long* gt_classes,
    float* gt_weights, float* gt_scores, float* gt_boxes, float* loss_weights, int* counters,
    void* stream);
/* label_and_sample_proposals with caller-supplied pseudo-gt seeds (roi_heads.py:263-370) -- the second half of
 * jtsm_refine_assign for seed sets it does not build itself: get_pgt_mist (roi_heads_oicr.py:549-591: top 15% per
 * class + class-agnostic NMS), thresholded / top-k > 1 get_pgt_top_k, or annotated boxes.
 * pgt_* [total seeds] DEVICE, seeds of image n at [seed_offsets[n], seed_offsets[n+1]) (int32 DEVICE);
 * max_seeds_per_image bounds the shared-memory staging (<= 5000).  Outputs as jtsm_refine_assign. */
JTSM_API int jtsm_refine_match(const float* pgt_boxes, const long long* pgt_classes, const float* pgt_scores,
                               const float* pgt_weights, const int* seed_offsets, int max_seeds_per_image,
                               const float* proposals, const int* seg_offsets, int num_images, int num_rois,
                               int num_classes, float iou_t0, float iou_t1, int label0, int label1, int label2,
                               int num_thresholds, long long* matched_idx, float* matched_iou,
                               long long* gt_classes, float* gt_weights, float* gt_scores, float* gt_boxes,
                               float* loss_weights, int* counters, void* stream);
/* jtsm_refine_match with the number of seeds per image read from DEVICE memory, so that seed sets whose size is data
 * dependent (thresholded get_pgt_top_k, roi_heads_oicr.py:754-760; the NMS survivors of get_pgt_mist, :573-581) need no
 * host read and the branch can sit inside a CUDA graph: image n owns the padded slot [seed_offsets[n],
 * seed_offsets[n+1]) of pgt_*, of which the first min(seed_counts[n], slot size) entries are seeds.
 * seed_counts int32 [num_images] DEVICE (NULL: every slot is full == jtsm_refine_match); max_seeds_per_image bounds the
 * SLOT size. */
JTSM_API int jtsm_refine_match_counted(const float* pgt_boxes, const long long* pgt_classes, const float* pgt_scores,
                                       const float* pgt_weights, const int* seed_offsets, const int* seed_counts,
                                       int max_seeds_per_image, const float* proposals, const int* seg_offsets,
                                       int num_images, int num_rois, int num_classes, float iou_t0, float iou_t1,
                                       int label0, int label1, int label2, int num_thresholds, long long* matched_idx,
                                       float* matched_iou, long long* gt_classes, float* gt_weights, float* gt_scores,
                                       float* gt_boxes, float* loss_weights, int* counters, void* stream);
JTSM_API size_t jtsm_refine_workspace_bytes(void);
/* Weighted softmax cross-entropy, forward + backward + probabilities in one launch
 * (OICROutputs.softmax_cross_entropy_loss + _log_accuracy, fast_rcnn_oicr.py:252-298; predict_probs :744-762):
 *   loss = loss_scale * sum_r w_r CE(logits_r, gt_r; ignore -1) / *valid_count
 *   grad_logits = (*grad_loss) * d loss / d logits.   probs / grad_logits / acc_counters may be NULL.
 * acc_counters int32[4]: #(argmax == gt), #fg, #(fg and argmax == gt), #(fg and argmax == background). */
JTSM_API int jtsm_refine_ce(const float* logits, int num_rois, int num_cols,
                            const long long* gt_classes, const float* loss_weights,
                            const int* valid_count, const float* grad_loss, float loss_scale,
                            float* probs, float* grad_logits, float* loss, int* acc_counters,
                            void* workspace, size_t workspace_bytes, void* stream);
/* Allocation-free variant of jtsm_refine_ce.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same arguments and results as jtsm_refine_ce without the workspace: the
 * kernel's scratch (grid-sum ticket + per-CTA loss partials) is a file-scope `static __device__` array of 8 KB
 * that every GPU loading the library holds until exit.  Calls no cudaMalloc / cudaFree / allocating routine.
 * Limit: num_rois <= JTSM_REFINE_CE_STATIC_MAX_ROIS (also returned by jtsm_refine_ce_static_max_rois());
 * larger -> JTSM_ERR_STATIC_LIMIT ("size exceeds the static scratch limit").  All calls on one device share the
 * array and it keeps the last call's contents (the call clears what must start at zero): stream-ordered calls are
 * safe, calls that OVERLAP on the GPU (different streams) race -- serialise them. */
#define JTSM_REFINE_CE_STATIC_MAX_ROIS 262144
JTSM_API int jtsm_refine_ce_static_max_rois(void);
JTSM_API int jtsm_refine_ce_static(const float* logits, int num_rois, int num_cols,
                                   const long long* gt_classes, const float* loss_weights,
                                   const int* valid_count, const float* grad_loss, float loss_scale,
                                   float* probs, float* grad_logits, float* loss, int* acc_counters,
                                   void* stream);
/* out = (accumulate ? out : 0) + scale * softmax(logits, dim=1)   (predict_probs_K, fast_rcnn_oicr.py:764-783) */
JTSM_API int jtsm_softmax_rows(const float* logits, int num_rois, int num_cols, float scale,
                               int accumulate, float* out, void* stream);
/* Box regression loss of a refinement branch (OICROutputs.box_reg_loss, fast_rcnn_oicr.py:300-380):
 * smooth-L1 (beta; < 1e-5 = L1) between pred_deltas[fg rows, gt-class columns] and
 * Box2BoxTransform.get_deltas(proposals, gt_boxes) (box_regression.py:38-71), times loss_weights[r]
 * when given ("smooth_l1_weighted"), summed, / num_rois.  grad_deltas dense [R, boxes_per_row*4] or NULL. */
JTSM_API int jtsm_refine_box_reg(const float* pred_deltas, int boxes_per_row,
                                 const float* proposals, const float* gt_boxes,
                                 const long long* gt_classes, const float* loss_weights,
                                 int num_rois, int num_classes, float wx, float wy, float ww,
                                 float wh, float smooth_l1_beta, const float* grad_loss,
                                 float loss_scale, float* loss, float* grad_deltas,
                                 void* workspace, size_t workspace_bytes, void* stream);
/* Box2BoxTransform.apply_deltas (box_regression.py:73-109): out [R, boxes_per_row, 4];
 * deltas are multiplied by delta_scale first (1.0, or 1/Kr for predict_boxes_K). */
JTSM_API int jtsm_apply_deltas(const float* deltas, const float* boxes, int num_rois,
                               int boxes_per_row, float wx, float wy, float ww, float wh,
                               float scale_clamp, float delta_scale, float* out, void* stream);
/* pairwise_iou(boxes1 [n1,4], boxes2 [n2,4]) -> out [n1, n2]  (detectron2/structures/boxes.py:369-391) */
JTSM_API int jtsm_pairwise_iou(const float* boxes1, int n1, const float* boxes2, int n2,
                               float* out, void* stream);

/* ------------------------------------------------------------------------
 * batched_nms (detectron2/layers/nms.py:10-30 -> torchvision.ops.nms per class):
 * greedy NMS inside each class (idxs int64 or NULL = one class), suppress IoU > iou_threshold,
 * candidates visited in descending score (ties: lower index first).
 * keep int64 [num_boxes]: first *num_keep entries = kept indices in descending-score order,
 * rest = -1.  num_keep int64 [1] DEVICE.  Never synchronises.
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_nms_workspace_bytes(int num_boxes);
JTSM_API int jtsm_batched_nms(const float* boxes, const float* scores, const long long* idxs,
                              int num_boxes, float iou_threshold, long long* keep,
                              long long* num_keep, void* workspace, size_t workspace_bytes,
                              void* stream);

/* ------------------------------------------------------------------------
 * RoIPool -> fc6 fused on the tensor cores (inference: no argmax is produced).  Replaces the chain
 *   box_pooler -> * (objectness_logits + 1) -> flatten -> fc6 (+ ReLU)
 * of projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:391-403 and box_head.py:59-68,90 without materialising the
 * pooled [R, C, P, P] tensor: the pooled values are produced as bf16 A-operand tiles in shared memory and consumed by
 * tcgen05.mma (fp32 accumulation in TMEM); the weight tiles arrive by TMA.
 * feat      [batch, height, width, channels] bf16, CHANNELS-LAST (a staged copy of the NCHW map), channels % 64 == 0
 * weight    [out_features, P*P*channels] bf16, BIN-MAJOR columns: weight[n, bin*channels + c] = fc6.weight[n, c*P*P + bin]
 *           (a one-time re-layout of the checkpoint), out_features % 512 == 0
 * out       [num_rois, out_features] fp32 = relu?(bias + pooled_bf16 x weight^T), pooled_bf16 = bf16(max * row_scale)
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_pool_fc6_workspace_bytes(int batch, int num_rois, int pooled_h, int pooled_w);
JTSM_API int jtsm_pool_fc6_forward(const void* feat_nhwc_bf16, const float* rois, int batch, int channels,
                                   int height, int width, int num_rois, float spatial_scale, int pooled_h,
                                   int pooled_w, const float* row_scale, const void* weight_binmajor_bf16,
                                   const float* bias, int out_features, int relu, float* out,
                                   void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------
 * fast_rcnn_inference_single_image on the device
 * (projects/WSL/wsl/modeling/roi_heads/fast_rcnn_oicr.py:95-163 == fast_rcnn_wsddn.py:94-162; the stock twin
 * detectron2/modeling/roi_heads/fast_rcnn.py is what GeneralizedRCNNWithTTAAVG calls at test_time_augmentation_avg.py:393):
 * rows with a non-finite box or score are dropped, the background column (last) is ignored, boxes are clipped to
 * (image_h, image_w), candidates are the (row, class) pairs with score > score_thresh in row-major (`nonzero`) order,
 * per-class NMS (IoU > nms_thresh suppresses), the topk best (topk < 0: all) are returned in descending score.
 * boxes  [num_rois, boxes_per_row, 4] (boxes_per_row = 1 or num_classes), scores [num_rois, num_classes + 1].
 * Outputs, `capacity` rows each (capacity >= topk, or >= num_rois * num_classes when topk < 0): det_boxes [cap, 4],
 * det_scores [cap], det_classes int64 (pred_classes), det_rows int64 (filter_inds[:, 0]: index among the VALID rows,
 * what the reference returns as the kept indices), det_pred_inds int64 (pred_inds: the original row);
 * rows >= *num_det are zero / -1.  num_det int64 [1] DEVICE.  Never synchronises, CUDA-graph capturable.
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_fast_rcnn_inference_workspace_bytes(int num_rois, int num_classes);
JTSM_API int jtsm_fast_rcnn_inference(const float* boxes, int boxes_per_row, const float* scores,
                                      int num_rois, int num_classes, float image_h, float image_w,
                                      float score_thresh, float nms_thresh, int topk, int capacity,
                                      float* det_boxes, float* det_scores, long long* det_classes,
                                      long long* det_rows, long long* det_pred_inds, long long* num_det,
                                      void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------
 * Test-time-augmentation average.
 * GeneralizedRCNNWithTTAAVG._get_augmented_boxes
 * (projects/WSL/wsl/modeling/test_time_augmentation_avg.py:364-390): the boxes of every augmented
 * pass are mapped back to the original image with `tfm.inverse().apply_box` (a host numpy round trip
 * per augmentation in the reference), then boxes and scores are averaged over the augmentations.
 *
 * A transform list is a sequence of at most JTSM_TTA_MAX_OPS stages, applied in order; every stage
 * follows fvcore's Transform.apply_box: the four corners go through apply_coords, the result is the
 * (min, max) box of the mapped corners.  Stage kinds:
 *   JTSM_TTA_NOOP   NoOpTransform                 (coords unchanged; still takes min / max)
 *   JTSM_TTA_HFLIP  HFlipTransform(width = a)     x -> a - x
 *   JTSM_TTA_SCALE  ResizeTransform               x -> x * a, y -> y * b  (a = float(new_w / w), b = float(new_h / h);
 *                                                  detectron2/data/transforms/transform.py:149-152)
 *   JTSM_TTA_END    terminates the list early
 * op_type / op_a / op_b are HOST arrays [num_aug * JTSM_TTA_MAX_OPS] (copied into the launch parameters).
 * ------------------------------------------------------------------------ */
#define JTSM_TTA_MAX_OPS 4
#define JTSM_TTA_MAX_AUG 32
#define JTSM_TTA_END 0
#define JTSM_TTA_NOOP 1
#define JTSM_TTA_HFLIP 2
#define JTSM_TTA_SCALE 3
/* boxes / scores: HOST arrays of num_aug DEVICE pointers, each [num_rois, boxes_per_row, 4] /
 * [num_rois, score_cols] fp32 contiguous (the per-augmentation outputs stay where inference left them).
 * mean_boxes [num_rois, boxes_per_row, 4], mean_scores [num_rois, score_cols]: sum over the augmentations in
 * list order, divided by num_aug (what torch.mean(dim=0) computes on the CPU).  scores / mean_scores may be
 * NULL (boxes only) and so may boxes / mean_boxes (scores only).  One launch. */
JTSM_API int jtsm_tta_merge(const void* const* boxes, const void* const* scores, int num_aug,
                            int num_rois, int boxes_per_row, int score_cols, const int* op_type,
                            const float* op_a, const float* op_b, float* mean_boxes,
                            float* mean_scores, void* stream);
/* transform_proposals (test_time_augmentation_avg.py:27-70): boxes [n, 4] -> transform list (forward) ->
 * Boxes.clip to (image_h, image_w) (detectron2/structures/boxes.py:195-211) -> out_boxes [n, 4] and
 * keep [n] uint8 = Boxes.nonempty(threshold = min_box_size) (boxes.py:213-225: width > t and height > t).
 * One transform list (op arrays of JTSM_TTA_MAX_OPS entries, HOST). */
JTSM_API int jtsm_tta_transform_boxes(const float* boxes, int num_boxes, const int* op_type,
                                      const float* op_a, const float* op_b, float image_h,
                                      float image_w, float min_box_size, float* out_boxes,
                                      unsigned char* keep, void* stream);

/* ------------------------------------------------------------------------
 * Superpixel-set IoU + Matcher: pairwise_iou_sp
 * (projects/WSL/wsl/modeling/roi_heads/roi_heads_jtsm.py:49-73) as called by
 * label_and_sample_proposals_wsl (:1627-1634), followed by Matcher.__call__
 * (detectron2/modeling/matcher.py:61-103, no low-quality matches).
 * oh1 [n1, S] int32 (targets' superpixel flags), oh2 [n2, S] int32 (proposals'), cnt_sp [S] int64
 * (pixels per superpixel) -> iou [n1, n2] fp32 = float(inter) / float(area1 + area2 - inter) with exact
 * integer sums (0 / 0 = NaN as in the reference).  Optional (both or none): matched_idx [n2] int64 =
 * first arg-max over the targets, matched_label [n2] int8 by the thresholds (as jtsm_refine_match).
 * n1 == 0: matched_idx = 0, matched_label = label0.  S <= 2844 (shared-memory staging).
 * ------------------------------------------------------------------------ */
JTSM_API int jtsm_pairwise_iou_sp(const int* oh1, int n1, const int* oh2, int n2,
                                  const long long* cnt_sp, int S, float* iou, float iou_t0,
                                  float iou_t1, int label0, int label1, int label2,
                                  int num_thresholds, long long* matched_idx,
                                  signed char* matched_label, void* stream);

/* ------------------------------------------------------------------------
 * PCL loss: wsl/layers/pcl_loss.py:8-88 over
 * projects/WSL/wsl/layers/csrc/pcl_loss/pcl_loss_cpu.cpp:8-115 (the reference copies the probabilities to
 * the host and runs this CPU code on every forward and backward).  All arrays fp32 as in the reference:
 * probs [R, C] (column 0 = background), labels / cls_loss_weights / gt_assignment [R],
 * pc_labels / pc_probs / pc_count / img_cls_loss_weights [P], im_labels [C].
 * forward: output [C] per-class sums, loss [1] = sum(output) / R (may be NULL).
 * backward: grad_probs [R, C], already divided by R; the upstream gradient is not applied (the reference
 * ignores it, pcl_loss_cpu.cpp:59-115).
 * ------------------------------------------------------------------------ */
JTSM_API int jtsm_pcl_loss_forward(const float* probs, const float* labels,
                                   const float* cls_loss_weights, const float* pc_labels,
                                   const float* pc_probs, const float* img_cls_loss_weights,
                                   const float* im_labels, int num_rois, int channels,
                                   int num_positive, float* output, float* loss, void* stream);
JTSM_API int jtsm_pcl_loss_backward(const float* probs, const float* labels,
                                    const float* cls_loss_weights, const float* gt_assignment,
                                    const float* pc_labels, const float* pc_probs,
                                    const float* pc_count, const float* img_cls_loss_weights,
                                    const float* im_labels, int num_rois, int channels,
                                    int num_positive, float* grad_probs, void* stream);

/* Proposal clusters of PCL (wsl/modeling/roi_heads/third_party/pcl.py:150-200 `_get_proposal_clusters`; numpy on the
 * host in the reference): rois [R,4], centers [G,4] (16-byte aligned), center_labels / center_scores [G] (class column
 * ids 1..K as floats, scores), cls_prob [R, num_cols].  Outputs fp32 as the reference hands them to pcl_loss:
 * labels / cls_loss_weights / gt_assignment [R] (IoU first-maximum; < fg_thresh: label 0, assignment -1;
 * < bg_thresh: weight 0), pc_labels / pc_probs / pc_count / img_cls_loss_weights [G].  G <= 1024.
 * The cluster CENTRES (`_get_graph_centers`: sklearn KMeans + greedy graph cover) stay host code. */
JTSM_API int jtsm_pcl_proposal_clusters(const float* rois, const float* centers, const float* center_labels,
                                        const float* center_scores, const float* cls_prob, int num_rois,
                                        int num_centers, int num_cols, float fg_thresh, float bg_thresh, float* labels,
                                        float* cls_loss_weights, float* gt_assignment, float* pc_labels,
                                        float* pc_probs, float* pc_count, float* img_cls_loss_weights, void* stream);

/* ------------------------------------------------------------------------
 * ROIMerge (the CMIL head's clique merge): wsl/layers/roi_merge.py:10-49 over
 * projects/WSL/wsl/layers/csrc/ROIMerge/ROIMerge_cpu.cpp:33-232 (forward), :234-287 (backward).  The reference op is
 * CPU only (ROIMerge.h:48-53), its caller copies scores + the R x R IoU matrix to the host every iteration
 * (fast_rcnn_wsddn.py:621-667).  All pointers DEVICE.
 *   scores [R] objectness (sum of the MIL scores), iou [R, R] pairwise IoU of the proposals, C / D [R, K] fp32.
 *   lambda: the merge threshold of this iteration (getlambda(), ROIMerge_cpu.cpp:11-17 -- a host scalar);
 *   top_k <= 256 (reference 200), window <= 64 (reference 40).
 * outputs: I int32 [R] cluster id of every RoI; IC int32 [R] members per id (0 past num_id);
 *   MC / MD [R, K]: rows < num_id = mean of the members' rows (the reference's running sum, bit for bit), rows past
 *   num_id zero (the reference returns [num_id, K]: slice with stats[0]);
 *   stats int32 [4] = num_id, num_top_id, max_clique, min_clique (ROIMerge_cpu.cpp:183-196).
 * Equal scores are ranked lowest index first (the reference's std::sort leaves ties unspecified).
 * backward: GC[n] = GMC[I[n]] / IC[I[n]] (same for D).
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_roi_merge_workspace_bytes(int num_rois);
JTSM_API int jtsm_roi_merge_forward(const float* scores, const float* iou, const float* C, const float* D,
                                    int num_rois, int num_classes, float lambda, int top_k, int window, float* MC,
                                    float* MD, int* I, int* IC, int* stats, void* workspace, size_t workspace_bytes,
                                    void* stream);
JTSM_API int jtsm_roi_merge_backward(const float* GMC, const float* GMD, const int* I, const int* IC, int num_rois,
                                     int num_classes, float* GC, float* GD, void* stream);

/* ------------------------------------------------------------------------
 * ROILabel (CMIL / UWSOD refinement labels): wsl/layers/roi_label.py:10-60 over
 * projects/WSL/wsl/layers/csrc/ROILabel/ROILabel_cpu.cpp:16-194 (CPU only, ROILabel.h:29-34; callers
 * roi_heads_cmil.py:419-431, roi_heads_uwsod.py).  All pointers DEVICE.
 *   S [R, score_stride] scores (column c = class c; extra columns ignored), U [R, R] pairwise IoU,
 *   L [K] image labels (== 1 switches a class on), CW [K] per-class weights or NULL (then the seed's score),
 *   perm int32 [R]: the order in which the RoIs claim the capped foreground / background slots, or NULL = index order.
 *   The reference draws it with std::random_shuffle seeded by time(0) (ROILabel_cpu.cpp:108-117); here it is an input
 *   so that the op is a function of its arguments (the module draws it with torch.randperm on the device).
 *   fg: IoU >= fg_thresh while at most num_pos were accepted before (num_pos + 1 in total, as the reference's
 *   `num_pos <= num_pos_`); else bg: bg_thresh_lo <= IoU < bg_thresh_hi while at most num_neg; else weight 0.
 * outputs: RL int32 [R] (class of the matched seed, K for background), RW [R];
 *   stats float [4] = accepted fg, accepted bg, sum of fg weights, sum of bg weights of this call (the reference's
 *   display accumulators P[10..13]).  At most 256 seeds (classes on x top_k) are used.
 * ------------------------------------------------------------------------ */
JTSM_API size_t jtsm_roi_label_workspace_bytes(int num_rois);
JTSM_API int jtsm_roi_label_forward(const float* S, int score_stride, const float* U, const float* L, const float* CW,
                                    const int* perm, int num_rois, int num_classes, float fg_thresh,
                                    float bg_thresh_hi, float bg_thresh_lo, int num_pos, int num_neg, int top_k, int* RL,
                                    float* RW, float* stats, void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* JTSM_B200_H_ */
