// Test-time-augmentation average on the device.
//
// Reference: GeneralizedRCNNWithTTAAVG._get_augmented_boxes, projects/WSL/wsl/modeling/test_time_augmentation_avg.py:
// 364-390 -- per augmentation a D2H copy, fvcore's numpy `tfm.inverse().apply_box`, an H2D copy, then torch.cat +
// torch.mean over the augmentations for boxes and scores.  Here: ONE launch that reads every augmentation's boxes and
// scores where inference left them (pointer table in the launch parameters), applies the inverse stage list in
// registers and writes the two means.  HBM bound: (A + 1) x R x (4K + K + 1) x 4 bytes, each read / written once.
//
// Arithmetic follows the host code exactly so the result is bit-equal to the reference on the CPU:
//   Transform.apply_box (fvcore): corners (x0,y0),(x1,y0),(x0,y1),(x1,y1) -> apply_coords -> min / max per axis;
//   HFlipTransform.apply_coords: x = width - x;  ResizeTransform.apply_coords (detectron2/data/transforms/
//   transform.py:149-152): x * float32(new_w / w), y * float32(new_h / h) in fp32;  torch.mean(dim=0) on the CPU:
//   sequential sum in list order, then a true division by A.
#include "common.cuh"

namespace jtsm {

struct TtaOps {
  int type[JTSM_TTA_MAX_OPS];
  float a[JTSM_TTA_MAX_OPS];
  float b[JTSM_TTA_MAX_OPS];
};

struct TtaMergeParams {
  const float4* boxes[JTSM_TTA_MAX_AUG];
  const float* scores[JTSM_TTA_MAX_AUG];
  TtaOps ops[JTSM_TTA_MAX_AUG];
};

// One stage list applied to one box, each stage = fvcore Transform.apply_box.
__device__ __forceinline__ float4 tta_apply_ops(float4 b, const TtaOps& o) {
#pragma unroll
  for (int s = 0; s < JTSM_TTA_MAX_OPS; ++s) {
    const int t = o.type[s];
    if (t == JTSM_TTA_END) break;
    float x0 = b.x, y0 = b.y, x1 = b.z, y1 = b.w;
    if (t == JTSM_TTA_HFLIP) {
      x0 = __fsub_rn(o.a[s], x0);
      x1 = __fsub_rn(o.a[s], x1);
    } else if (t == JTSM_TTA_SCALE) {
      x0 = __fmul_rn(x0, o.a[s]);
      x1 = __fmul_rn(x1, o.a[s]);
      y0 = __fmul_rn(y0, o.b[s]);
      y1 = __fmul_rn(y1, o.b[s]);
    }
    b = make_float4(fminf(x0, x1), fminf(y0, y1), fmaxf(x0, x1), fmaxf(y0, y1));
  }
  return b;
}

// CTAs [0, box_ctas): one thread per box; CTAs [box_ctas, ...): one thread per score element.
__global__ void __launch_bounds__(256) tta_merge_kernel(const __grid_constant__ TtaMergeParams p, int num_aug,
                                                        long long num_boxes, long long num_scores, int box_ctas,
                                                        float4* __restrict__ mean_boxes,
                                                        float* __restrict__ mean_scores) {
  const float fa = (float)num_aug;
  if ((int)blockIdx.x < box_ctas) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= num_boxes) return;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    // four independent 16-byte loads in flight per thread; the sum itself stays in list order
    for (int a0 = 0; a0 < num_aug; a0 += 4) {
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (a0 + j < num_aug) v[j] = __ldg(p.boxes[a0 + j] + i);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (a0 + j >= num_aug) break;
        float4 b = tta_apply_ops(v[j], p.ops[a0 + j]);
        acc.x = __fadd_rn(acc.x, b.x);
        acc.y = __fadd_rn(acc.y, b.y);
        acc.z = __fadd_rn(acc.z, b.z);
        acc.w = __fadd_rn(acc.w, b.w);
      }
    }
    mean_boxes[i] = make_float4(__fdiv_rn(acc.x, fa), __fdiv_rn(acc.y, fa), __fdiv_rn(acc.z, fa),
                                __fdiv_rn(acc.w, fa));
  } else {
    const long long i = (long long)(blockIdx.x - box_ctas) * 256 + threadIdx.x;
    if (i >= num_scores) return;
    float acc = 0.f;
    for (int a0 = 0; a0 < num_aug; a0 += 4) {
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (a0 + j < num_aug) v[j] = __ldg(p.scores[a0 + j] + i);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (a0 + j < num_aug) acc = __fadd_rn(acc, v[j]);
    }
    mean_scores[i] = __fdiv_rn(acc, fa);
  }
}

__global__ void __launch_bounds__(256) tta_transform_boxes_kernel(const float4* __restrict__ boxes, int n,
                                                                  const __grid_constant__ TtaOps ops, float image_h,
                                                                  float image_w, float min_box_size,
                                                                  float4* __restrict__ out,
                                                                  unsigned char* __restrict__ keep) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  float4 b = tta_apply_ops(__ldg(boxes + i), ops);
  // Boxes.clip (boxes.py:195-211): clamp(min=0, max=w / h)
  b.x = fminf(fmaxf(b.x, 0.f), image_w);
  b.y = fminf(fmaxf(b.y, 0.f), image_h);
  b.z = fminf(fmaxf(b.z, 0.f), image_w);
  b.w = fminf(fmaxf(b.w, 0.f), image_h);
  out[i] = b;
  // Boxes.nonempty (boxes.py:213-225)
  keep[i] = (__fsub_rn(b.z, b.x) > min_box_size) && (__fsub_rn(b.w, b.y) > min_box_size);
}

static bool fill_ops(TtaOps* o, const int* type, const float* a, const float* b) {
  for (int s = 0; s < JTSM_TTA_MAX_OPS; ++s) {
    if (type[s] < JTSM_TTA_END || type[s] > JTSM_TTA_SCALE) return false;
    o->type[s] = type[s];
    o->a[s] = a[s];
    o->b[s] = b[s];
  }
  return true;
}

}  // namespace jtsm

using namespace jtsm;

extern "C" int jtsm_tta_merge(const void* const* boxes, const void* const* scores, int num_aug, int num_rois,
                              int boxes_per_row, int score_cols, const int* op_type, const float* op_a,
                              const float* op_b, float* mean_boxes, float* mean_scores, void* stream) {
  if (num_aug <= 0 || num_aug > JTSM_TTA_MAX_AUG || num_rois < 0 || boxes_per_row < 0 || score_cols < 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois == 0) return JTSM_OK;
  const bool do_boxes = boxes && mean_boxes && boxes_per_row > 0;
  const bool do_scores = scores && mean_scores && score_cols > 0;
  if (!do_boxes && !do_scores) return JTSM_ERR_INVALID_ARGUMENT;
  if (do_boxes && (!op_type || !op_a || !op_b)) return JTSM_ERR_INVALID_ARGUMENT;
  TtaMergeParams p;
  for (int a = 0; a < JTSM_TTA_MAX_AUG; ++a) {
    p.boxes[a] = nullptr;
    p.scores[a] = nullptr;
    for (int s = 0; s < JTSM_TTA_MAX_OPS; ++s) p.ops[a].type[s] = JTSM_TTA_END, p.ops[a].a[s] = p.ops[a].b[s] = 0.f;
  }
  for (int a = 0; a < num_aug; ++a) {
    if (do_boxes) {
      if (!boxes[a] || ((uintptr_t)boxes[a] & 15)) return JTSM_ERR_INVALID_ARGUMENT;
      p.boxes[a] = (const float4*)boxes[a];
      if (!fill_ops(&p.ops[a], op_type + a * JTSM_TTA_MAX_OPS, op_a + a * JTSM_TTA_MAX_OPS,
                    op_b + a * JTSM_TTA_MAX_OPS))
        return JTSM_ERR_INVALID_ARGUMENT;
    }
    if (do_scores) {
      if (!scores[a]) return JTSM_ERR_INVALID_ARGUMENT;
      p.scores[a] = (const float*)scores[a];
    }
  }
  if (do_boxes && ((uintptr_t)mean_boxes & 15)) return JTSM_ERR_INVALID_ARGUMENT;
  const long long nb = do_boxes ? (long long)num_rois * boxes_per_row : 0;
  const long long ns = do_scores ? (long long)num_rois * score_cols : 0;
  const long long bc = (nb + 255) / 256, sc = (ns + 255) / 256;
  if (bc + sc > 0x7fffffffLL) return JTSM_ERR_INVALID_ARGUMENT;
  tta_merge_kernel<<<(unsigned)(bc + sc), 256, 0, (cudaStream_t)stream>>>(p, num_aug, nb, ns, (int)bc,
                                                                          (float4*)mean_boxes, mean_scores);
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_tta_transform_boxes(const float* boxes, int num_boxes, const int* op_type, const float* op_a,
                                        const float* op_b, float image_h, float image_w, float min_box_size,
                                        float* out_boxes, unsigned char* keep, void* stream) {
  if (num_boxes < 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_boxes == 0) return JTSM_OK;
  if (!boxes || !out_boxes || !keep || !op_type || !op_a || !op_b) return JTSM_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)boxes & 15) || ((uintptr_t)out_boxes & 15)) return JTSM_ERR_INVALID_ARGUMENT;
  TtaOps o;
  if (!fill_ops(&o, op_type, op_a, op_b)) return JTSM_ERR_INVALID_ARGUMENT;
  tta_transform_boxes_kernel<<<(num_boxes + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      (const float4*)boxes, num_boxes, o, image_h, image_w, min_box_size, (float4*)out_boxes, keep);
  count_launch();
  return cuda_status();
}
