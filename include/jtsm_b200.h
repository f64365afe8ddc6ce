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

#ifdef __cplusplus
}
#endif
#endif /* JTSM_B200_H_ */
