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
