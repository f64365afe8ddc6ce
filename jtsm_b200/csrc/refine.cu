// OICR-style refinement for sm_100a: pseudo ground truth (top-scoring proposal per image-level
// class), proposal x seed IoU label assignment, weighted cross-entropy (+ weighted smooth-L1 box
// regression), softmax for the next branch, and box decoding.
//
// Reference (eager PyTorch, ~50 tiny kernels and 3-7 host syncs per branch and image):
//   get_pgt_top_k            projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:675-811
//   label_and_sample_proposals  projects/WSL/wsl/modeling/roi_heads/roi_heads.py:263-370
//   _sample_proposals        projects/WSL/wsl/modeling/roi_heads/roi_heads.py:222-254
//   pairwise_iou             detectron2/structures/boxes.py:345-391
//   Matcher.__call__         detectron2/modeling/matcher.py:61-103
//   OICROutputs (weights, CE, box reg)  projects/WSL/wsl/modeling/roi_heads/fast_rcnn_oicr.py:243-247,282-298,300-380
//   Box2BoxTransform         detectron2/modeling/box_regression.py:38-109
//
// Bit-exactness notes: IoU uses the reference's operation order with explicit round-to-nearest
// intrinsics (no FMA contraction, IEEE division); argmax ties resolve to the lowest index
// (what torch.max(dim=0) / CPU topk do).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace jtsm {

constexpr int kAssignThreads = 256;
constexpr int kAssignCluster = 8;  // CTAs per image

struct MatcherParams {  // detectron2/modeling/matcher.py:36-59
  float t0, t1;         // thresholds (t1 unused when n == 1)
  int l0, l1, l2;       // labels for [-inf,t0), [t0,t1), [t1,inf)   (n == 1: l0, l1)
  int n;
};

__device__ __forceinline__ float box_area(float x1, float y1, float x2, float y2) {
  return __fmul_rn(__fsub_rn(x2, x1), __fsub_rn(y2, y1));  // Boxes.area, boxes.py:184-193
}

// iou(gt, prop) exactly as pairwise_iou(boxes1=gt, boxes2=prop): boxes.py:361-390
__device__ __forceinline__ float iou_ref(const float4 g, float area_g, const float4 p, float area_p) {
  float w = __fsub_rn(fminf(g.z, p.z), fmaxf(g.x, p.x));
  float h = __fsub_rn(fminf(g.w, p.w), fmaxf(g.y, p.y));
  w = fmaxf(w, 0.f);
  h = fmaxf(h, 0.f);
  const float inter = __fmul_rn(w, h);
  if (!(inter > 0.f)) return 0.f;
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(area_g, area_p), inter));
}

// One thread-block CLUSTER of kAssignCluster CTAs per image (a single CTA was latency bound at ~14 us for 2000
// proposals); proposals are strided over all threads of the cluster.
//  phase 1: for every image-level class g, argmax over the image's proposals of prev_scores[:, cls_g] (lowest
//           index on ties): per-CTA candidates, exchanged through distributed shared memory, every CTA then holds
//           the seed boxes / classes / scores / weights;   (skipped when the caller supplies the seeds)
//  phase 2: every proposal: IoU against the G seeds, first max, Matcher thresholds -> labels.
__device__ __forceinline__ bool seed_better(float ov, int oi, float bv, int bi) {
  return oi != 0x7fffffff && (bi == 0x7fffffff || ov > bv || (ov == bv && oi < bi));
}

__global__ void __launch_bounds__(kAssignThreads)
refine_assign_kernel(const float* __restrict__ prev_scores, int score_stride,
                     const float* __restrict__ prev_boxes, int boxes_per_row,
                     const float* __restrict__ proposals, const int* __restrict__ seg,
                     int num_classes, const int* __restrict__ cls_off,
                     const long long* __restrict__ cls_ids, const float* __restrict__ img_scores,
                     MatcherParams mp, float* __restrict__ pgt_boxes,
                     long long* __restrict__ pgt_classes, float* __restrict__ pgt_scores,
                     float* __restrict__ pgt_weights, long long* __restrict__ matched_idx,
                     float* __restrict__ matched_iou, long long* __restrict__ gt_classes,
                     float* __restrict__ gt_weights, float* __restrict__ gt_scores,
                     float* __restrict__ gt_boxes, float* __restrict__ loss_weights,
                     int* __restrict__ counters /* [N*3] fg,bg,ignore ; then [1] valid; zeroed by the caller */,
                     int num_images, const int* __restrict__ seed_counts /* optional: device-side seeds per image */) {
  extern __shared__ float4 seeds[];  // [G] boxes, then float area[G], weight[G], score[G], int cls[G], cand_v[G], cand_i[G]
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank(), nrank = (int)cluster.num_blocks();
  const int n = blockIdx.x / nrank;
  const int r0 = seg[n], r1 = seg[n + 1];
  const int g0 = cls_off[n];
  // caller-supplied seeds may come padded: slot [cls_off[n], cls_off[n+1]) with seed_counts[n] of them in use
  const int G = seed_counts ? min(max(seed_counts[n], 0), cls_off[n + 1] - g0) : cls_off[n + 1] - g0;
  float* s_area = (float*)(seeds + G);
  float* s_weight = s_area + G;
  float* s_score = s_weight + G;
  int* s_cls = (int*)(s_score + G);
  float* cand_v = (float*)(s_cls + G);
  int* cand_i = (int*)(cand_v + G);
  __shared__ float red_v[kAssignThreads / 32];
  __shared__ int red_i[kAssignThreads / 32];
  __shared__ int s_cnt[4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tid = rank * kAssignThreads + threadIdx.x, nthr = nrank * kAssignThreads;
  if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0;
  const bool has_gt = (G > 0) && (r1 > r0);  // top_k = min(num_pred, 1)
  const bool given = prev_scores == nullptr;  // seeds supplied by the caller (jtsm_refine_match)
  if (given) {
    for (int g = threadIdx.x; g < G; g += kAssignThreads) {
      const float4 b = *(const float4*)(pgt_boxes + (size_t)(g0 + g) * 4);
      seeds[g] = b;
      s_area[g] = box_area(b.x, b.y, b.z, b.w);
      s_weight[g] = pgt_weights[g0 + g];
      s_score[g] = pgt_scores[g0 + g];
      s_cls[g] = (int)pgt_classes[g0 + g];
    }
  } else {
    // per-CTA candidate of every class
    for (int g = 0; g < G && has_gt; ++g) {
      const int cls = (int)cls_ids[g0 + g];
      float bv = -INFINITY;
      int bi = 0x7fffffff;
      for (int r = r0 + tid; r < r1; r += nthr) {
        const float v = prev_scores[(size_t)r * score_stride + cls];
        if (v > bv || bi == 0x7fffffff) {
          bv = v;
          bi = r;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (seed_better(ov, oi, bv, bi)) bv = ov, bi = oi;
      }
      if (lane == 0) red_v[warp] = bv, red_i[warp] = bi;
      __syncthreads();
      if (warp == 0) {
        bv = lane < kAssignThreads / 32 ? red_v[lane] : -INFINITY;
        bi = lane < kAssignThreads / 32 ? red_i[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (seed_better(ov, oi, bv, bi)) bv = ov, bi = oi;
        }
        if (lane == 0) cand_v[g] = bv, cand_i[g] = bi;
      }
      __syncthreads();
    }
    cluster.sync();  // every CTA's candidates are visible cluster-wide
    for (int g = threadIdx.x; g < G; g += kAssignThreads) {
      if (has_gt) {
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int j = 0; j < nrank; ++j) {  // fixed order: deterministic
          const float ov = cluster.map_shared_rank(cand_v, j)[g];
          const int oi = cluster.map_shared_rank(cand_i, j)[g];
          if (seed_better(ov, oi, bv, bi)) bv = ov, bi = oi;
        }
        const int cls = (int)cls_ids[g0 + g];
        const int col = boxes_per_row > 1 ? cls : 0;
        const float4 b = *(const float4*)(prev_boxes + ((size_t)bi * boxes_per_row + col) * 4);
        const float wgt = img_scores[(size_t)n * num_classes + cls];
        seeds[g] = b;
        s_area[g] = box_area(b.x, b.y, b.z, b.w);
        s_weight[g] = wgt;
        s_score[g] = bv;
        s_cls[g] = cls;
        if (rank == 0) {
          *(float4*)(pgt_boxes + (size_t)(g0 + g) * 4) = b;
          pgt_classes[g0 + g] = cls;
          pgt_scores[g0 + g] = bv;
          pgt_weights[g0 + g] = wgt;
        }
      } else if (rank == 0) {  // image without proposals: its seed slots hold zeros
        *(float4*)(pgt_boxes + (size_t)(g0 + g) * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        pgt_classes[g0 + g] = 0;
        pgt_scores[g0 + g] = 0.f;
        pgt_weights[g0 + g] = 0.f;
      }
    }
  }
  cluster.sync();  // seeds complete in every CTA; nobody reads remote candidates after this point

  int c_fg = 0, c_bg = 0, c_ig = 0, c_valid = 0;
  for (int r = r0 + tid; r < r1; r += nthr) {
    const float4 p = *(const float4*)(proposals + (size_t)r * 4);
    long long cls_out;
    float w_out = 0.f, sc_out = 0.f, best = 0.f;
    int bi = 0;
    float4 gb = p;  // placeholder when the image has no seeds (fast_rcnn_oicr.py:228-231)
    if (has_gt) {
      const float ap = box_area(p.x, p.y, p.z, p.w);
      best = -1.f;
      for (int g = 0; g < G; ++g) {
        const float v = iou_ref(seeds[g], s_area[g], p, ap);
        if (v > best) {  // first max wins (matcher.py:92)
          best = v;
          bi = g;
        }
      }
      int label;
      if (mp.n == 1) label = best >= mp.t0 ? mp.l1 : mp.l0;
      else label = best >= mp.t1 ? mp.l2 : (best >= mp.t0 ? mp.l1 : mp.l0);
      cls_out = s_cls[bi];
      if (label == 0) cls_out = num_classes;
      if (label == -1) cls_out = -1;
      w_out = s_weight[bi];
      sc_out = s_score[bi];
      gb = seeds[bi];
    } else {
      cls_out = num_classes;
    }
    matched_idx[r] = bi;
    if (matched_iou) matched_iou[r] = best;
    gt_classes[r] = cls_out;
    gt_weights[r] = w_out;
    gt_scores[r] = sc_out;
    *(float4*)(gt_boxes + (size_t)r * 4) = gb;
    const float lw = cls_out == -1 ? 0.f : w_out;  // fast_rcnn_oicr.py:243-244
    loss_weights[r] = lw;
    c_ig += cls_out == -1;
    c_bg += cls_out == num_classes;
    c_fg += (cls_out != -1 && cls_out != num_classes);
    c_valid += lw > 1e-12f;  // fast_rcnn_oicr.py:246-247
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c_fg += __shfl_xor_sync(0xffffffffu, c_fg, o);
    c_bg += __shfl_xor_sync(0xffffffffu, c_bg, o);
    c_ig += __shfl_xor_sync(0xffffffffu, c_ig, o);
    c_valid += __shfl_xor_sync(0xffffffffu, c_valid, o);
  }
  if (lane == 0) {
    atomicAdd(&s_cnt[0], c_fg);
    atomicAdd(&s_cnt[1], c_bg);
    atomicAdd(&s_cnt[2], c_ig);
    atomicAdd(&s_cnt[3], c_valid);
  }
  __syncthreads();
  // integer sums: order independent
  if (threadIdx.x < 3) atomicAdd(counters + n * 3 + threadIdx.x, s_cnt[threadIdx.x]);
  if (threadIdx.x == 3) atomicAdd(counters + num_images * 3, s_cnt[3]);
}

// Deterministic grid-wide float sum: per-CTA partial -> the last CTA to arrive adds the partials
// with warp 0 (lane l takes partials l, l+32, ... in order, then a fixed shuffle tree).
// Called by every thread of the CTA; `cta_value` is read from thread 0.
__device__ __forceinline__ void grid_sum_finalize(float cta_value, float* partials, unsigned int* ticket,
                                                  float* out, float divisor, float multiplier) {
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = cta_value;
    __threadfence();
    s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last || threadIdx.x >= 32) return;
  __threadfence();
  float t = 0.f;
  for (unsigned int i = threadIdx.x; i < gridDim.x; i += 32) t += ((volatile float*)partials)[i];
  t = warp_sum(t);
  if (threadIdx.x == 0) {
    *out = t / divisor * multiplier;
    *ticket = 0u;
  }
}

__device__ __forceinline__ float block_sum_256(float v, float* s_red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x == 0)
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_red[w];
  return t;  // valid in thread 0
}

// Weighted softmax cross-entropy forward + backward + probabilities, one warp per row.
//   loss = loss_scale * sum_r w_r * CE(logits_r, gt_r; ignore -1) / valid      (fast_rcnn_oicr.py:293-298)
//   dlogits_rj = grad_loss * loss_scale * w_r / valid * (p_rj - [j == gt_r])
template <int KT>
__global__ void __launch_bounds__(256)
refine_ce_kernel(const float* __restrict__ logits, int Kc, const long long* __restrict__ gt,
                 const float* __restrict__ w, const int* __restrict__ valid_count,
                 const float* __restrict__ grad_loss, int num_rois, float loss_scale,
                 float* __restrict__ probs, float* __restrict__ grad_logits,
                 float* __restrict__ loss, int* __restrict__ acc, float* partials,
                 unsigned int* ticket) {
  __shared__ float s_red[8];
  __shared__ int s_acc[4];
  if (threadIdx.x < 4) s_acc[threadIdx.x] = 0;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float valid = (float)(*valid_count);
  const float gscale = (grad_loss ? *grad_loss : 1.f) * loss_scale;
  float part = 0.f;
  int a_acc = 0, a_fg = 0, a_fgacc = 0, a_fn = 0;
  for (int r = blockIdx.x * 8 + warp; r < num_rois; r += gridDim.x * 8) {
    float x[KT];
    float m = -INFINITY;
    int am = 0x7fffffff;
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      x[t] = k < Kc ? logits[(size_t)r * Kc + k] : -INFINITY;
      if (k < Kc && (x[t] > m || am == 0x7fffffff)) {
        m = x[t];
        am = k;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o);
      const int oa = __shfl_xor_sync(0xffffffffu, am, o);
      if (oa != 0x7fffffff && (am == 0x7fffffff || om > m || (om == m && oa < am))) {
        m = om;
        am = oa;
      }
    }
    float s = 0.f;
#pragma unroll
    for (int t = 0; t < KT; ++t)
      if (lane + 32 * t < Kc) s += expf(x[t] - m);
    s = warp_sum(s);
    const float inv = 1.f / s;
    const float lse = m + logf(s);
    const long long g = gt[r];
    const float wr = w[r];
    const bool ign = (g < 0);
    const float coef = ign ? 0.f : gscale * wr / valid;
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < Kc) {
        const float p = expf(x[t] - m) * inv;
        if (probs) probs[(size_t)r * Kc + k] = p;
        if (grad_logits) grad_logits[(size_t)r * Kc + k] = coef * (p - (k == (int)g ? 1.f : 0.f));
        if (!ign && k == (int)g) part += wr * (lse - x[t]);  // exactly one lane per row
      }
    }
    if (lane == 0) {  // _log_accuracy, fast_rcnn_oicr.py:252-280
      const int bg = Kc - 1;
      const bool fg = (g >= 0 && g < bg);
      a_acc += (am == (int)g);
      a_fg += fg;
      a_fgacc += fg && (am == (int)g);
      a_fn += fg && (am == bg);
    }
  }
  if (lane == 0 && acc) {
    atomicAdd(&s_acc[0], a_acc);
    atomicAdd(&s_acc[1], a_fg);
    atomicAdd(&s_acc[2], a_fgacc);
    atomicAdd(&s_acc[3], a_fn);
  }
  const float cta = block_sum_256(part, s_red);
  grid_sum_finalize(cta, partials, ticket, loss, valid, loss_scale);
  if (acc && threadIdx.x < 4) atomicAdd(acc + threadIdx.x, s_acc[threadIdx.x]);
}

// softmax rows, optionally accumulated:  out = (accumulate ? out : 0) + scale * softmax(logits)
// (predict_probs / predict_probs_K, fast_rcnn_oicr.py:744-783)
template <int KT>
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float* __restrict__ logits, int Kc, int num_rois, float scale,
                    int accumulate, float* __restrict__ out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = blockIdx.x * 8 + warp; r < num_rois; r += gridDim.x * 8) {
    float x[KT];
    float m = -INFINITY;
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      x[t] = k < Kc ? logits[(size_t)r * Kc + k] : -INFINITY;
      m = fmaxf(m, x[t]);
    }
    m = warp_max(m);
    float s = 0.f;
#pragma unroll
    for (int t = 0; t < KT; ++t)
      if (lane + 32 * t < Kc) s += expf(x[t] - m);
    s = warp_sum(s);
    const float inv = 1.f / s;
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < Kc) {
        const float p = expf(x[t] - m) * inv * scale;
        float* o = out + (size_t)r * Kc + k;
        *o = accumulate ? *o + p : p;
      }
    }
  }
}

struct BoxWeights {
  float wx, wy, ww, wh;
};

// Box2BoxTransform.get_deltas for one coordinate j of (src -> target); box_regression.py:52-66
__device__ __forceinline__ float delta_target(const float4 s, const float4 t, const BoxWeights bw, int j) {
  const float sw = s.z - s.x, sh = s.w - s.y;
  const float tw = t.z - t.x, th = t.w - t.y;
  switch (j) {
    case 0: return bw.wx * ((t.x + 0.5f * tw) - (s.x + 0.5f * sw)) / sw;
    case 1: return bw.wy * ((t.y + 0.5f * th) - (s.y + 0.5f * sh)) / sh;
    case 2: return bw.ww * logf(tw / sw);
    default: return bw.wh * logf(th / sh);
  }
}

// Box regression loss (smooth-L1 / L1, optionally weighted by the proposal weight) over the
// gt-class columns of foreground rows, / R; dense gradient.  fast_rcnn_oicr.py:300-380.
__global__ void __launch_bounds__(256)
refine_box_reg_kernel(const float* __restrict__ deltas, int boxes_per_row,
                      const float* __restrict__ proposals, const float* __restrict__ gt_boxes,
                      const long long* __restrict__ gt, const float* __restrict__ w, int num_rois,
                      int num_classes, BoxWeights bw, float beta, const float* __restrict__ grad_loss,
                      float loss_scale, float* __restrict__ loss, float* __restrict__ grad_deltas,
                      float* partials, unsigned int* ticket) {
  __shared__ float s_red[8];
  const int cols = boxes_per_row * 4;
  const long long total = (long long)num_rois * cols;
  const float gscale = (grad_loss ? *grad_loss : 1.f) * loss_scale / (float)num_rois;
  float part = 0.f;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(e / cols);
    const int c = (int)(e - (long long)r * cols);
    const long long g = gt[r];
    float grad = 0.f;
    if (g >= 0 && g < num_classes) {
      const int base = boxes_per_row > 1 ? 4 * (int)g : 0;
      if (c >= base && c < base + 4) {
        const float4 s = *(const float4*)(proposals + (size_t)r * 4);
        const float4 t = *(const float4*)(gt_boxes + (size_t)r * 4);
        const float tgt = delta_target(s, t, bw, c - base);
        const float d = deltas[e] - tgt;
        const float ad = fabsf(d);
        float l, dl;
        if (beta < 1e-5f) {  // fvcore smooth_l1_loss: pure L1 below beta 1e-5
          l = ad;
          dl = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        } else if (ad < beta) {
          l = 0.5f * d * d / beta;
          dl = d / beta;
        } else {
          l = ad - 0.5f * beta;
          dl = d > 0.f ? 1.f : -1.f;
        }
        const float wr = w ? w[r] : 1.f;
        part += l * wr;
        grad = dl * wr * gscale;
      }
    }
    if (grad_deltas) grad_deltas[e] = grad;
  }
  const float cta = block_sum_256(part, s_red);
  grid_sum_finalize(cta, partials, ticket, loss, (float)num_rois, loss_scale);
}

// Box2BoxTransform.apply_deltas; box_regression.py:73-109.  One thread per (row, class) box.
__global__ void apply_deltas_kernel(const float* __restrict__ deltas, const float* __restrict__ boxes,
                                    int num_rois, int boxes_per_row, BoxWeights bw, float scale_clamp,
                                    float delta_scale, float* __restrict__ out) {
  const long long total = (long long)num_rois * boxes_per_row;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(e / boxes_per_row);
    const float4 b = *(const float4*)(boxes + (size_t)r * 4);
    float4 d = *(const float4*)(deltas + e * 4);
    d.x *= delta_scale; d.y *= delta_scale; d.z *= delta_scale; d.w *= delta_scale;
    const float wd = b.z - b.x, ht = b.w - b.y;
    const float cx = b.x + 0.5f * wd, cy = b.y + 0.5f * ht;
    const float dx = d.x / bw.wx, dy = d.y / bw.wy;
    const float dw = fminf(d.z / bw.ww, scale_clamp), dh = fminf(d.w / bw.wh, scale_clamp);
    const float pcx = __fadd_rn(__fmul_rn(dx, wd), cx), pcy = __fadd_rn(__fmul_rn(dy, ht), cy);
    const float pw = expf(dw) * wd, ph = expf(dh) * ht;
    float4 o;
    o.x = __fsub_rn(pcx, __fmul_rn(0.5f, pw));
    o.y = __fsub_rn(pcy, __fmul_rn(0.5f, ph));
    o.z = __fadd_rn(pcx, __fmul_rn(0.5f, pw));
    o.w = __fadd_rn(pcy, __fmul_rn(0.5f, ph));
    *(float4*)(out + e * 4) = o;
  }
}

// pairwise_iou(boxes1 [n1,4], boxes2 [n2,4]) -> [n1, n2]; boxes.py:369-391.
// Tile of 32 boxes1 rows staged in shared memory per CTA; threads stride over boxes2.
__global__ void __launch_bounds__(256)
pairwise_iou_kernel(const float* __restrict__ b1, int n1, const float* __restrict__ b2, int n2,
                    float* __restrict__ out) {
  __shared__ float4 s_b[32];
  __shared__ float s_a[32];
  const int i0 = blockIdx.y * 32;
  if (threadIdx.x < 32 && i0 + threadIdx.x < n1) {
    const float4 b = *(const float4*)(b1 + (size_t)(i0 + threadIdx.x) * 4);
    s_b[threadIdx.x] = b;
    s_a[threadIdx.x] = box_area(b.x, b.y, b.z, b.w);
  }
  __syncthreads();
  const int rows = min(32, n1 - i0);
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n2; j += gridDim.x * blockDim.x) {
    const float4 p = *(const float4*)(b2 + (size_t)j * 4);
    const float ap = box_area(p.x, p.y, p.z, p.w);
    for (int i = 0; i < rows; ++i) out[(size_t)(i0 + i) * n2 + j] = iou_ref(s_b[i], s_a[i], p, ap);
  }
}

}  // namespace jtsm

using namespace jtsm;

namespace {
struct SumWs {
  unsigned int* ticket;
  float* partials;
};
constexpr int kMaxSumCtas = 148 * 8;
SumWs carve_sum_ws(void* workspace) {
  char* p = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  return {(unsigned int*)p, (float*)(p + 16)};
}
}  // namespace

namespace {
template <typename... Args>
cudaError_t launch_assign(int num_images, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(num_images * kAssignCluster));
  cfg.blockDim = dim3(kAssignThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kAssignCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, refine_assign_kernel, args...);
}
}  // namespace

extern "C" size_t jtsm_refine_workspace_bytes(void) { return 16 + kMaxSumCtas * 4 + 512; }

extern "C" int jtsm_refine_assign(
    const float* prev_scores, int score_stride, const float* prev_boxes, int boxes_per_row,
    const float* proposals, const int* seg_offsets, int num_images, int num_rois, int num_classes,
    const int* img_cls_offsets, const long long* img_cls_ids, int max_classes_per_image,
    const float* img_scores, float iou_t0, float iou_t1, int label0, int label1, int label2,
    int num_thresholds, float* pgt_boxes, long long* pgt_classes, float* pgt_scores,
    float* pgt_weights, long long* matched_idx, float* matched_iou, long long* gt_classes,
    float* gt_weights, float* gt_scores, float* gt_boxes, float* loss_weights, int* counters,
    void* stream) {
  if (num_images < 0 || num_rois < 0 || num_classes <= 0 || score_stride < num_classes ||
      (boxes_per_row != 1 && boxes_per_row != num_classes) || num_thresholds < 1 ||
      num_thresholds > 2 || max_classes_per_image < 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (!counters) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(counters, 0, ((size_t)num_images * 3 + 1) * sizeof(int), st);
  if (num_images == 0) return JTSM_OK;
  if (!seg_offsets || !img_cls_offsets || !matched_idx || !gt_classes || !gt_weights ||
      !gt_scores || !gt_boxes || !loss_weights)
    return JTSM_ERR_INVALID_ARGUMENT;
  const size_t smem = (size_t)max_classes_per_image * (16 + 6 * 4) + 16;
  if (smem > 200 * 1024) return JTSM_ERR_UNSUPPORTED;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(refine_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  MatcherParams mp{iou_t0, iou_t1, label0, label1, label2, num_thresholds};
  launch_assign(num_images, smem, st, prev_scores, score_stride, prev_boxes, boxes_per_row, proposals, seg_offsets,
                num_classes, img_cls_offsets, img_cls_ids, img_scores, mp, pgt_boxes, pgt_classes, pgt_scores,
                pgt_weights, matched_idx, matched_iou, gt_classes, gt_weights, gt_scores, gt_boxes, loss_weights,
                counters, num_images, (const int*)nullptr);
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_refine_match_counted(const float* pgt_boxes, const long long* pgt_classes,
                                         const float* pgt_scores, const float* pgt_weights, const int* seed_offsets,
                                         const int* seed_counts, int max_seeds_per_image, const float* proposals,
                                         const int* seg_offsets, int num_images, int num_rois, int num_classes,
                                         float iou_t0, float iou_t1, int label0, int label1, int label2,
                                         int num_thresholds, long long* matched_idx, float* matched_iou,
                                         long long* gt_classes, float* gt_weights, float* gt_scores, float* gt_boxes,
                                         float* loss_weights, int* counters, void* stream) {
  if (num_images < 0 || num_rois < 0 || num_classes <= 0 || num_thresholds < 1 || num_thresholds > 2 ||
      max_seeds_per_image < 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (!counters) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(counters, 0, ((size_t)num_images * 3 + 1) * sizeof(int), st);
  if (num_images == 0) return JTSM_OK;
  if (!seg_offsets || !seed_offsets || !matched_idx || !gt_classes || !gt_weights || !gt_scores || !gt_boxes ||
      !loss_weights)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (max_seeds_per_image > 0 && (!pgt_boxes || !pgt_classes || !pgt_scores || !pgt_weights))
    return JTSM_ERR_INVALID_ARGUMENT;
  const size_t smem = (size_t)max_seeds_per_image * (16 + 6 * 4) + 16;
  if (smem > 200 * 1024) return JTSM_ERR_UNSUPPORTED;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(refine_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  MatcherParams mp{iou_t0, iou_t1, label0, label1, label2, num_thresholds};
  launch_assign(num_images, smem, st, (const float*)nullptr, 0, (const float*)nullptr, 1, proposals, seg_offsets,
                num_classes, seed_offsets, (const long long*)nullptr, (const float*)nullptr, mp,
                const_cast<float*>(pgt_boxes), const_cast<long long*>(pgt_classes), const_cast<float*>(pgt_scores),
                const_cast<float*>(pgt_weights), matched_idx, matched_iou, gt_classes, gt_weights, gt_scores, gt_boxes,
                loss_weights, counters, num_images, seed_counts);
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_refine_match(const float* pgt_boxes, const long long* pgt_classes, const float* pgt_scores,
                                 const float* pgt_weights, const int* seed_offsets, int max_seeds_per_image,
                                 const float* proposals, const int* seg_offsets, int num_images, int num_rois,
                                 int num_classes, float iou_t0, float iou_t1, int label0, int label1, int label2,
                                 int num_thresholds, long long* matched_idx, float* matched_iou,
                                 long long* gt_classes, float* gt_weights, float* gt_scores, float* gt_boxes,
                                 float* loss_weights, int* counters, void* stream) {
  return jtsm_refine_match_counted(pgt_boxes, pgt_classes, pgt_scores, pgt_weights, seed_offsets, nullptr,
                                   max_seeds_per_image, proposals, seg_offsets, num_images, num_rois, num_classes, iou_t0,
                                   iou_t1, label0, label1, label2, num_thresholds, matched_idx, matched_iou, gt_classes,
                                   gt_weights, gt_scores, gt_boxes, loss_weights, counters, stream);
}

extern "C" int jtsm_refine_ce(const float* logits, int num_rois, int num_cols,
                              const long long* gt_classes, const float* loss_weights,
                              const int* valid_count, const float* grad_loss, float loss_scale,
                              float* probs, float* grad_logits, float* loss, int* acc_counters,
                              void* workspace, size_t workspace_bytes, void* stream) {
  if (num_rois < 0 || num_cols <= 0 || num_cols > 256) return JTSM_ERR_INVALID_ARGUMENT;
  if (!loss || !workspace || !valid_count) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_refine_workspace_bytes()) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  if (num_rois > 0 && (!logits || !gt_classes || !loss_weights)) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  SumWs ws = carve_sum_ws(workspace);
  cudaMemsetAsync(ws.ticket, 0, 4, st);
  if (acc_counters) cudaMemsetAsync(acc_counters, 0, 4 * sizeof(int), st);
  int grid = (num_rois + 7) / 8;
  grid = grid < 1 ? 1 : (grid > kMaxSumCtas ? kMaxSumCtas : grid);
  JTSM_DISPATCH_KT(num_cols, KT, {
    refine_ce_kernel<KT><<<grid, 256, 0, st>>>(logits, num_cols, gt_classes, loss_weights, valid_count,
                                                grad_loss, num_rois, loss_scale, probs, grad_logits, loss,
                                                acc_counters, ws.partials, ws.ticket);
  });
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_softmax_rows(const float* logits, int num_rois, int num_cols, float scale,
                                 int accumulate, float* out, void* stream) {
  if (num_rois < 0 || num_cols <= 0 || num_cols > 256) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois == 0) return JTSM_OK;
  if (!logits || !out) return JTSM_ERR_INVALID_ARGUMENT;
  int grid = (num_rois + 7) / 8;
  grid = grid > kMaxSumCtas ? kMaxSumCtas : grid;
  JTSM_DISPATCH_KT(num_cols, KT, {
    softmax_rows_kernel<KT><<<grid, 256, 0, (cudaStream_t)stream>>>(logits, num_cols, num_rois, scale,
                                                                     accumulate, out);
  });
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_refine_box_reg(const float* pred_deltas, int boxes_per_row,
                                   const float* proposals, const float* gt_boxes,
                                   const long long* gt_classes, const float* loss_weights,
                                   int num_rois, int num_classes, float wx, float wy, float ww,
                                   float wh, float smooth_l1_beta, const float* grad_loss,
                                   float loss_scale, float* loss, float* grad_deltas,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  if (num_rois < 0 || num_classes <= 0 || (boxes_per_row != 1 && boxes_per_row != num_classes))
    return JTSM_ERR_INVALID_ARGUMENT;
  if (!loss || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_refine_workspace_bytes()) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  if (num_rois > 0 && (!pred_deltas || !proposals || !gt_boxes || !gt_classes))
    return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  SumWs ws = carve_sum_ws(workspace);
  cudaMemsetAsync(ws.ticket, 0, 4, st);
  const long long total = (long long)num_rois * boxes_per_row * 4;
  long long grid = (total + 255) / 256;
  grid = grid < 1 ? 1 : (grid > kMaxSumCtas ? kMaxSumCtas : grid);
  refine_box_reg_kernel<<<(int)grid, 256, 0, st>>>(pred_deltas, boxes_per_row, proposals, gt_boxes,
                                                    gt_classes, loss_weights, num_rois, num_classes,
                                                    BoxWeights{wx, wy, ww, wh}, smooth_l1_beta, grad_loss,
                                                    loss_scale, loss, grad_deltas, ws.partials, ws.ticket);
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_apply_deltas(const float* deltas, const float* boxes, int num_rois,
                                 int boxes_per_row, float wx, float wy, float ww, float wh,
                                 float scale_clamp, float delta_scale, float* out, void* stream) {
  if (num_rois < 0 || boxes_per_row <= 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois == 0) return JTSM_OK;
  if (!deltas || !boxes || !out) return JTSM_ERR_INVALID_ARGUMENT;
  const long long total = (long long)num_rois * boxes_per_row;
  long long grid = (total + 255) / 256;
  grid = grid > kMaxSumCtas ? kMaxSumCtas : grid;
  apply_deltas_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>(
      deltas, boxes, num_rois, boxes_per_row, BoxWeights{wx, wy, ww, wh}, scale_clamp, delta_scale, out);
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_pairwise_iou(const float* boxes1, int n1, const float* boxes2, int n2,
                                 float* out, void* stream) {
  if (n1 < 0 || n2 < 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (n1 == 0 || n2 == 0) return JTSM_OK;
  if (!boxes1 || !boxes2 || !out) return JTSM_ERR_INVALID_ARGUMENT;
  dim3 grid((n2 + 255) / 256, (n1 + 31) / 32);
  if (grid.x > 4096) grid.x = 4096;
  pairwise_iou_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(boxes1, n1, boxes2, n2, out);
  count_launch();
  return cuda_status();
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of jtsm_refine_ce (no counterpart in the reference; added
// for callers that cannot allocate: CUDA-graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  The kernel's scratch -- the grid-sum ticket and the per-CTA loss partials -- lives in a file-scope
// device array instead of a caller-provided workspace.  Every GPU that loads the library holds the 8 KB until exit.
// All calls on a device share the array: calls that overlap on the GPU (different streams) race; stream-ordered
// calls are safe (the ticket is cleared at the start of every call, the partials are overwritten before they are read).
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr size_t kStaticScratchBytes = 8192;  // 16 (ticket) + kMaxSumCtas * 4 (partials) = 4752, rounded up
static_assert(16 + (size_t)kMaxSumCtas * 4 <= kStaticScratchBytes, "static scratch too small for the grid-sum");
static __device__ __align__(256) unsigned char g_refine_ce_scratch[kStaticScratchBytes];

// In host code `g_refine_ce_scratch` is a host-side stand-in, not a device address: the device address is looked up
// once per device (with that device current) and cached by device ordinal.
int static_scratch_ptr(void** out) {
  constexpr int kMaxDev = 64;
  static std::atomic<void*> cached[kMaxDev];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return JTSM_ERR_CUDA_BASE + (int)e;
  if (dev < 0 || dev >= kMaxDev) return JTSM_ERR_UNSUPPORTED;
  void* p = cached[dev].load(std::memory_order_acquire);
  if (!p) {
    e = cudaGetSymbolAddress(&p, g_refine_ce_scratch);
    if (e != cudaSuccess || !p) return JTSM_ERR_CUDA_BASE + (int)(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
    cached[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return JTSM_OK;
}
}  // namespace

extern "C" int jtsm_refine_ce_static_max_rois(void) { return JTSM_REFINE_CE_STATIC_MAX_ROIS; }

extern "C" int jtsm_refine_ce_static(const float* logits, int num_rois, int num_cols, const long long* gt_classes,
                                     const float* loss_weights, const int* valid_count, const float* grad_loss,
                                     float loss_scale, float* probs, float* grad_logits, float* loss,
                                     int* acc_counters, void* stream) {
  if (num_rois < 0 || num_cols <= 0 || num_cols > 256) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois > JTSM_REFINE_CE_STATIC_MAX_ROIS) return JTSM_ERR_STATIC_LIMIT;  // beyond the stated maximum
  if (!loss || !valid_count) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois > 0 && (!logits || !gt_classes || !loss_weights)) return JTSM_ERR_INVALID_ARGUMENT;
  void* scratch = nullptr;
  const int rc = static_scratch_ptr(&scratch);
  if (rc != JTSM_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* ticket = (unsigned int*)scratch;
  float* partials = (float*)((char*)scratch + 16);
  cudaMemsetAsync(ticket, 0, 4, st);  // the array keeps the last call's contents: the ticket must start at zero
  if (acc_counters) cudaMemsetAsync(acc_counters, 0, 4 * sizeof(int), st);
  int grid = (num_rois + 7) / 8;
  grid = grid < 1 ? 1 : (grid > kMaxSumCtas ? kMaxSumCtas : grid);
  JTSM_DISPATCH_KT(num_cols, KT, {
    refine_ce_kernel<KT><<<grid, 256, 0, st>>>(logits, num_cols, gt_classes, loss_weights, valid_count, grad_loss,
                                                num_rois, loss_scale, probs, grad_logits, loss, acc_counters, partials,
                                                ticket);
  });
  count_launch();
  return cuda_status();
}
