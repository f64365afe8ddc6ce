// fast_rcnn_inference_single_image on the device, for sm_100a.
//
// Reference: projects/WSL/wsl/modeling/roi_heads/fast_rcnn_oicr.py:95-163 (identical twin fast_rcnn_wsddn.py:94-162):
//   valid rows = all boxes and all scores of the row finite (rows are compacted when any is invalid);
//   drop the background column; Boxes.clip to the image (detectron2/structures/boxes.py:195-211);
//   candidates = (row, class) with score > score_thresh, in `nonzero` (row-major) order;
//   per-class NMS (detectron2/layers/nms.py:10-30), keep[:topk];
//   result: boxes, scores, classes, pred_inds (original row) and filter_inds[:, 0] (row among the VALID rows).
// The reference does this with ~15 eager kernels, a `nonzero` host sync and torchvision's NMS (device bitmask +
// host-side scan).  Here: 4 small kernels + one ordered stream compaction (CUB select over a counting iterator) in
// front of the NMS pipeline of nms.cu, whose box count is read on the device -- no host sync, graph-capturable.
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include "common.cuh"

namespace jtsm {

int nms_run(const float* boxes, const float* scores, const long long* idxs, int n, const int* num_valid,
            float iou_threshold, long long* keep, long long* num_keep, void* workspace, size_t workspace_bytes,
            cudaStream_t st);

// warp per row: valid[r] = every box coordinate and every score (background included) is finite
__global__ void infer_row_valid_kernel(const float* __restrict__ boxes, int box_cols, const float* __restrict__ scores,
                                       int score_cols, int num_rois, int* __restrict__ valid) {
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (r >= num_rois) return;
  bool ok = true;
  for (int j = lane; j < box_cols; j += 32) ok &= isfinite(boxes[(size_t)r * box_cols + j]);
  for (int j = lane; j < score_cols; j += 32) ok &= isfinite(scores[(size_t)r * score_cols + j]);
  ok = __all_sync(0xffffffffu, ok);
  if (lane == 0) valid[r] = ok ? 1 : 0;
}

// candidate predicate over flat (row, class) ids, row-major == torch.nonzero order
struct InferCand {
  const float* scores;
  const int* valid;
  int K;
  float thresh;
  __device__ __forceinline__ bool operator()(const int& i) const {
    const int r = i / K, k = i - r * K;
    return valid[r] && scores[(size_t)r * (K + 1) + k] > thresh;
  }
};

// candidate j < *num_cand: clipped box, score, class
__global__ void infer_gather_kernel(const int* __restrict__ cand, const int* __restrict__ num_cand,
                                    const float* __restrict__ boxes, int boxes_per_row,
                                    const float* __restrict__ scores, int K, float img_h, float img_w,
                                    float4* __restrict__ cboxes, float* __restrict__ cscores,
                                    long long* __restrict__ ccls, int capacity) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= capacity || j >= *num_cand) return;
  const int i = cand[j], r = i / K, k = i - r * K;
  const float4 b = *(const float4*)(boxes + ((size_t)r * boxes_per_row + (boxes_per_row == 1 ? 0 : k)) * 4);
  // Boxes.clip: x in [0, w], y in [0, h]
  cboxes[j] = make_float4(fminf(fmaxf(b.x, 0.f), img_w), fminf(fmaxf(b.y, 0.f), img_h), fminf(fmaxf(b.z, 0.f), img_w),
                          fminf(fmaxf(b.w, 0.f), img_h));
  cscores[j] = scores[(size_t)r * (K + 1) + k];
  ccls[j] = k;
}

// detections t < min(*num_keep, topk): gather through keep -> candidate -> (row, class)
__global__ void infer_emit_kernel(const long long* __restrict__ keep, const long long* __restrict__ num_keep, int topk,
                                  const int* __restrict__ cand, const float4* __restrict__ cboxes,
                                  const float* __restrict__ cscores, const int* __restrict__ vrow, int K,
                                  float4* __restrict__ det_boxes, float* __restrict__ det_scores,
                                  long long* __restrict__ det_classes, long long* __restrict__ det_rows,
                                  long long* __restrict__ det_pred_inds, long long* __restrict__ num_det, int cap) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  long long nk = *num_keep;
  if (topk >= 0 && nk > topk) nk = topk;
  if (nk > cap) nk = cap;
  if (t == 0) *num_det = nk;
  if (t >= cap) return;
  if (t >= nk) {  // padding rows: defined contents
    det_boxes[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    det_scores[t] = 0.f;
    det_classes[t] = -1;
    det_rows[t] = -1;
    det_pred_inds[t] = -1;
    return;
  }
  const int j = (int)keep[t], i = cand[j], r = i / K;
  det_boxes[t] = cboxes[j];
  det_scores[t] = cscores[j];
  det_classes[t] = i - r * K;
  det_rows[t] = vrow[r];  // filter_inds[:, 0]: index among the valid rows
  det_pred_inds[t] = r;   // pred_inds: the original row
}

struct InferWs {
  int *valid, *vrow, *cand, *num_cand;
  float4* cboxes;
  float* cscores;
  long long *ccls, *keep, *num_keep;
  void* cub_temp;
  size_t cub_bytes;
  void* nms_ws;
  size_t nms_bytes;
};

}  // namespace jtsm

using namespace jtsm;

extern "C" size_t jtsm_nms_workspace_bytes(int num_boxes);

namespace {
size_t infer_layout(int R, int K, void* base, InferWs* out) {
  char* p = (char*)(((uintptr_t)base + 255) & ~(uintptr_t)255);
  char* p0 = p;
  auto take = [&](size_t bytes) {
    char* q = p;
    p += align_up(bytes, 256);
    return q;
  };
  const size_t n = (size_t)R * K;
  InferWs w;
  w.valid = (int*)take((size_t)R * 4);
  w.vrow = (int*)take((size_t)R * 4);
  w.cand = (int*)take(n * 4);
  w.num_cand = (int*)take(16);
  w.cboxes = (float4*)take(n * 16);
  w.cscores = (float*)take(n * 4);
  w.ccls = (long long*)take(n * 8);
  w.keep = (long long*)take(n * 8);
  w.num_keep = (long long*)take(16);
  w.cub_bytes = n * 8 + (2u << 20);
  w.cub_temp = take(w.cub_bytes);
  w.nms_bytes = jtsm_nms_workspace_bytes((int)n);
  w.nms_ws = take(w.nms_bytes);
  if (out) *out = w;
  return (size_t)(p - p0) + 256;
}
}  // namespace

extern "C" size_t jtsm_fast_rcnn_inference_workspace_bytes(int num_rois, int num_classes) {
  if (num_rois < 0 || num_classes <= 0 || (long long)num_rois * num_classes > 0x3FFFFFFFLL) return 0;
  return infer_layout(num_rois, num_classes, nullptr, nullptr);
}

extern "C" int jtsm_fast_rcnn_inference(const float* boxes, int boxes_per_row, const float* scores, int num_rois,
                                        int num_classes, float image_h, float image_w, float score_thresh,
                                        float nms_thresh, int topk, int capacity, float* det_boxes,
                                        float* det_scores, long long* det_classes, long long* det_rows,
                                        long long* det_pred_inds, long long* num_det, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  if (num_rois < 0 || num_classes <= 0 || capacity < 0 || !num_det) return JTSM_ERR_INVALID_ARGUMENT;
  if (boxes_per_row != 1 && boxes_per_row != num_classes) return JTSM_ERR_INVALID_ARGUMENT;
  if ((long long)num_rois * num_classes > 0x3FFFFFFFLL) return JTSM_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  const int R = num_rois, K = num_classes, n = R * K;
  if (capacity > 0 && (!det_boxes || !det_scores || !det_classes || !det_rows || !det_pred_inds))
    return JTSM_ERR_INVALID_ARGUMENT;
  if (n == 0) {
    cudaMemsetAsync(num_det, 0, sizeof(long long), st);
    return cuda_status();
  }
  if (!boxes || !scores || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_fast_rcnn_inference_workspace_bytes(R, K)) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  InferWs w;
  infer_layout(R, K, workspace, &w);
  infer_row_valid_kernel<<<(R * 32 + 255) / 256, 256, 0, st>>>(boxes, boxes_per_row * 4, scores, K + 1, R, w.valid);
  size_t need = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, need, w.valid, w.vrow, R, st);
  if (need > w.cub_bytes) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cub::DeviceScan::ExclusiveSum(w.cub_temp, need, w.valid, w.vrow, R, st);
  thrust::counting_iterator<int> ids(0);
  InferCand pred{scores, w.valid, K, score_thresh};
  need = 0;
  cub::DeviceSelect::If(nullptr, need, ids, w.cand, w.num_cand, n, pred, st);
  if (need > w.cub_bytes) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cub::DeviceSelect::If(w.cub_temp, need, ids, w.cand, w.num_cand, n, pred, st);
  infer_gather_kernel<<<(n + 255) / 256, 256, 0, st>>>(w.cand, w.num_cand, boxes, boxes_per_row, scores, K, image_h,
                                                       image_w, w.cboxes, w.cscores, w.ccls, n);
  count_launch(4);
  int rc = nms_run((const float*)w.cboxes, w.cscores, w.ccls, n, w.num_cand, nms_thresh, w.keep, w.num_keep, w.nms_ws,
                   w.nms_bytes, st);
  if (rc != JTSM_OK) return rc;
  const int cap = capacity;
  const int blocks = cap > 0 ? (cap + 255) / 256 : 1;
  infer_emit_kernel<<<blocks, 256, 0, st>>>(w.keep, w.num_keep, topk, w.cand, w.cboxes, w.cscores, w.vrow, K,
                                            (float4*)det_boxes, det_scores, det_classes, det_rows, det_pred_inds, num_det,
                                            cap);
  count_launch();
  return cuda_status();
}
