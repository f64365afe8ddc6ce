// PCL loss on the device.
//
// Reference: wsl/layers/pcl_loss.py:8-88 copies the probabilities to the CPU on every call (forward AND backward) and
// runs projects/WSL/wsl/layers/csrc/pcl_loss/pcl_loss_cpu.cpp (its CUDA branch is compiled out, pcl_loss.h:64,99).
//   forward  (:8-57):  out[c] = 0 unless im_labels[c] != 0;
//                      c == 0: - sum_{i: labels[i] == 0} cls_loss_weights[i] * log(max(prob[i, 0], 1e-6))
//                      c  > 0: - sum_{j: pc_labels[j] == c} img_cls_loss_weights[j] * log(max(pc_probs[j], 1e-6))
//                      loss = sum_c out[c] / R                                          (pcl_loss.py:52)
//   backward (:59-115): grad[i, c] = 0 unless im_labels[c] != 0;
//                      c == 0 and labels[i] == 0: - cls_loss_weights[i] / max(prob[i, 0], 1e-5)
//                      c  > 0 and labels[i] == c: - img_cls_loss_weights[p] / max(pc_count[p] * pc_probs[p], 1e-5),
//                                                  p = gt_assignment[i];   then grad /= R   (pcl_loss.py:84)
//                      (the upstream gradient is not applied -- the reference ignores top_grad; kept.)
// Forward: one CTA, a warp per class, lane-strided partial sums combined in a fixed order (deterministic; the
// reference's serial sum differs in association -> tolerance parity, like every loss).  Backward: elementwise,
// round-to-nearest divisions in the reference's order -> bit-exact.
#include "common.cuh"

namespace jtsm {

__global__ void __launch_bounds__(1024) pcl_loss_fwd_kernel(const float* __restrict__ probs,
                                                            const float* __restrict__ labels,
                                                            const float* __restrict__ clw,
                                                            const float* __restrict__ pc_labels,
                                                            const float* __restrict__ pc_probs,
                                                            const float* __restrict__ iclw,
                                                            const float* __restrict__ im_labels, int R, int C, int P,
                                                            float* __restrict__ out, float* __restrict__ loss) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const float eps = 1e-6f;
  for (int c = warp; c < C; c += nwarp) {
    float acc = 0.f;
    if (im_labels[c] != 0.f) {
      if (c == 0) {
        for (int i = lane; i < R; i += 32)
          if (labels[i] == 0.f) acc -= clw[i] * logf(fmaxf(probs[(size_t)i * C], eps));
      } else {
        for (int j = lane; j < P; j += 32)
          if (pc_labels[j] == (float)c) acc -= iclw[j] * logf(fmaxf(pc_probs[j], eps));
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) out[c] = acc;
  }
  __syncthreads();
  if (threadIdx.x == 0 && loss) {
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += out[c];
    *loss = s / (float)R;
  }
}

__global__ void __launch_bounds__(256) pcl_loss_bwd_kernel(const float* __restrict__ probs,
                                                           const float* __restrict__ labels,
                                                           const float* __restrict__ clw,
                                                           const float* __restrict__ gt_assignment,
                                                           const float* __restrict__ pc_labels,
                                                           const float* __restrict__ pc_probs,
                                                           const float* __restrict__ pc_count,
                                                           const float* __restrict__ iclw,
                                                           const float* __restrict__ im_labels, int R, int C, int P,
                                                           float* __restrict__ grad) {
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)R * C) return;
  const int i = (int)(idx / C), c = (int)(idx - (long long)i * C);
  const float eps = 1e-5f;
  float g = 0.f;
  if (im_labels[c] != 0.f) {
    const float lab = labels[i];
    if (c == 0) {
      if (lab == 0.f) g = __fdiv_rn(-clw[i], fmaxf(probs[idx], eps));
    } else if (lab == (float)c) {
      const int p = (int)gt_assignment[i];
      if (p >= 0 && p < P) g = __fdiv_rn(-iclw[p], fmaxf(__fmul_rn(pc_count[p], pc_probs[p]), eps));
    }
  }
  grad[idx] = __fdiv_rn(g, (float)R);
}


// ---------------------------------------------------------------------------------------------------------------
// Proposal clusters (third_party/pcl.py:150-200 `_get_proposal_clusters`): every proposal joins the cluster centre of
// highest IoU (first maximum); below fg_thresh it is background (label 0, assignment -1), below bg_thresh its loss weight
// is 0; per centre: number of members, sum of their weights, mean of their probabilities in the centre's class column.
// The reference does this with numpy on the host after copying boxes and probabilities from the device.
constexpr int kPclMaxCenters = 1024;

__device__ __forceinline__ float pcl_iou(const float4 a, const float4 b) {  // detectron2 pairwise_iou, boxes.py:361-390
  const float area_a = __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(a.w, a.y));
  const float area_b = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
  const float w = fmaxf(__fsub_rn(fminf(a.z, b.z), fmaxf(a.x, b.x)), 0.f);
  const float h = fmaxf(__fsub_rn(fminf(a.w, b.w), fmaxf(a.y, b.y)), 0.f);
  const float inter = __fmul_rn(w, h);
  if (!(inter > 0.f)) return 0.f;
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(area_a, area_b), inter));
}

__global__ void __launch_bounds__(256) pcl_assign_kernel(const float4* __restrict__ rois, const float4* __restrict__ centers,
                                                         const float* __restrict__ center_labels,
                                                         const float* __restrict__ center_scores, int R, int G,
                                                         float fg_thresh, float bg_thresh, float* __restrict__ labels,
                                                         float* __restrict__ weights, float* __restrict__ assignment) {
  extern __shared__ float4 s_centers[];
  for (int g = threadIdx.x; g < G; g += blockDim.x) s_centers[g] = centers[g];
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R) return;
  const float4 b = rois[i];
  float best = -1.f;
  int bi = 0;
  for (int g = 0; g < G; ++g) {
    const float u = pcl_iou(b, s_centers[g]);
    if (u > best) {  // numpy argmax: first maximum
      best = u;
      bi = g;
    }
  }
  const bool bg = best < fg_thresh;
  labels[i] = bg ? 0.f : center_labels[bi];
  weights[i] = best < bg_thresh ? 0.f : center_scores[bi];
  assignment[i] = bg ? -1.f : (float)bi;
}

// One CTA per centre; fixed-order sums (lane-strided partials, shuffle tree, warp partials in order): deterministic.
__global__ void __launch_bounds__(256) pcl_cluster_stats_kernel(const float* __restrict__ assignment,
                                                                const float* __restrict__ weights,
                                                                const float* __restrict__ cls_prob, int C,
                                                                const float* __restrict__ center_labels, int R,
                                                                float* __restrict__ pc_labels, float* __restrict__ pc_probs,
                                                                float* __restrict__ pc_count,
                                                                float* __restrict__ img_cls_loss_weights) {
  __shared__ float s_w[8], s_p[8];
  __shared__ int s_n[8];
  const int g = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float lab = center_labels[g];
  const int col = (int)lab;
  float w = 0.f, p = 0.f;
  int n = 0;
  for (int i = threadIdx.x; i < R; i += blockDim.x) {
    if (assignment[i] == (float)g) {
      w += weights[i];
      p += (col >= 0 && col < C) ? cls_prob[(size_t)i * C + col] : 0.f;
      n++;
    }
  }
  for (int o = 16; o; o >>= 1) {
    w += __shfl_xor_sync(0xffffffffu, w, o);
    p += __shfl_xor_sync(0xffffffffu, p, o);
    n += __shfl_xor_sync(0xffffffffu, n, o);
  }
  if (lane == 0) {
    s_w[warp] = w;
    s_p[warp] = p;
    s_n[warp] = n;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ws = 0.f, ps = 0.f;
    int ns = 0;
    for (int k = 0; k < 8; ++k) {
      ws += s_w[k];
      ps += s_p[k];
      ns += s_n[k];
    }
    pc_labels[g] = lab;
    pc_count[g] = (float)ns;
    img_cls_loss_weights[g] = ws;
    pc_probs[g] = __fdiv_rn(ps, (float)ns);  // np.average of an empty set is NaN in the reference too (0 / 0)
  }
}

}  // namespace jtsm

using namespace jtsm;

extern "C" int jtsm_pcl_loss_forward(const float* probs, const float* labels, const float* cls_loss_weights,
                                     const float* pc_labels, const float* pc_probs,
                                     const float* img_cls_loss_weights, const float* im_labels, int num_rois,
                                     int channels, int num_positive, float* output, float* loss, void* stream) {
  if (num_rois <= 0 || channels <= 0 || num_positive < 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (!probs || !labels || !cls_loss_weights || !im_labels || !output) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_positive > 0 && (!pc_labels || !pc_probs || !img_cls_loss_weights)) return JTSM_ERR_INVALID_ARGUMENT;
  const int threads = std::min(1024, ((channels + 0) * 32 + 31) / 32 * 32);
  pcl_loss_fwd_kernel<<<1, std::max(32, threads), 0, (cudaStream_t)stream>>>(
      probs, labels, cls_loss_weights, pc_labels, pc_probs, img_cls_loss_weights, im_labels, num_rois, channels,
      num_positive, output, loss);
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_pcl_loss_backward(const float* probs, const float* labels, const float* cls_loss_weights,
                                      const float* gt_assignment, const float* pc_labels, const float* pc_probs,
                                      const float* pc_count, const float* img_cls_loss_weights,
                                      const float* im_labels, int num_rois, int channels, int num_positive,
                                      float* grad_probs, void* stream) {
  if (num_rois <= 0 || channels <= 0 || num_positive < 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (!probs || !labels || !cls_loss_weights || !gt_assignment || !im_labels || !grad_probs)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (num_positive > 0 && (!pc_labels || !pc_probs || !pc_count || !img_cls_loss_weights))
    return JTSM_ERR_INVALID_ARGUMENT;
  const long long n = (long long)num_rois * channels;
  pcl_loss_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      probs, labels, cls_loss_weights, gt_assignment, pc_labels, pc_probs, pc_count, img_cls_loss_weights, im_labels,
      num_rois, channels, num_positive, grad_probs);
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_pcl_proposal_clusters(const float* rois, const float* centers, const float* center_labels,
                                          const float* center_scores, const float* cls_prob, int num_rois,
                                          int num_centers, int num_cols, float fg_thresh, float bg_thresh, float* labels,
                                          float* cls_loss_weights, float* gt_assignment, float* pc_labels,
                                          float* pc_probs, float* pc_count, float* img_cls_loss_weights, void* stream) {
  if (num_rois < 0 || num_centers <= 0 || num_cols <= 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_centers > kPclMaxCenters) return JTSM_ERR_UNSUPPORTED;
  if (!centers || !center_labels || !center_scores || !pc_labels || !pc_probs || !pc_count || !img_cls_loss_weights)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois > 0 && (!rois || !cls_prob || !labels || !cls_loss_weights || !gt_assignment))
    return JTSM_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)rois | (uintptr_t)centers) % 16 != 0) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  if (num_rois > 0) {
    pcl_assign_kernel<<<(num_rois + 255) / 256, 256, (size_t)num_centers * sizeof(float4), st>>>(
        (const float4*)rois, (const float4*)centers, center_labels, center_scores, num_rois, num_centers, fg_thresh,
        bg_thresh, labels, cls_loss_weights, gt_assignment);
    count_launch();
  }
  pcl_cluster_stats_kernel<<<num_centers, 256, 0, st>>>(gt_assignment, cls_loss_weights, cls_prob, num_cols,
                                                        center_labels, num_rois, pc_labels, pc_probs, pc_count,
                                                        img_cls_loss_weights);
  count_launch();
  return cuda_status();
}
