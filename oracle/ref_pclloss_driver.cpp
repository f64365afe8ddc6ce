// extern "C" driver around the reference's OWN CPU PCL loss (projects/WSL/wsl/layers/csrc/pcl_loss/pcl_loss_cpu.cpp:
// 8-57 forward, :59-115 backward; the Python wrapper wsl/layers/pcl_loss.py:24-52 always runs this CPU code, its
// CUDA branch is disabled in pcl_loss.h:64,99), compiled in place by oracle/Makefile.ref against this image's ATen.
// TEST INFRASTRUCTURE ONLY: pins oracle/jtsm_oracle.c's PCL-loss restatement; never linked into the product.
#include <ATen/ATen.h>

#include "pcl_loss/pcl_loss.h"

static at::Tensor view(const float* p, long rows, long cols) {
  return at::from_blob(const_cast<float*>(p), {rows, cols}, at::kFloat);
}

extern "C" {

void ref_pcl_loss_forward(const float* probs, const float* labels, const float* cls_loss_weights,
                          const float* pc_labels, const float* pc_probs, const float* img_cls_loss_weights,
                          const float* im_labels, int batch, int channels, int num_positive, float* output) {
  at::Tensor a = view(probs, batch, channels), b = view(labels, 1, batch), c = view(cls_loss_weights, 1, batch),
             d = view(pc_labels, 1, num_positive), e = view(pc_probs, 1, num_positive),
             f = view(img_cls_loss_weights, 1, num_positive), g = view(im_labels, 1, channels),
             o = view(output, 1, channels);
  wsl::pcl_loss_forward_cpu(a, b, c, d, e, f, g, o);
}

void ref_pcl_loss_backward(const float* probs, const float* labels, const float* cls_loss_weights,
                           const float* gt_assignment, const float* pc_labels, const float* pc_probs,
                           const float* pc_count, const float* img_cls_loss_weights, const float* im_labels,
                           int batch, int channels, int num_positive, float* bottom_grad) {
  at::Tensor a = view(probs, batch, channels), b = view(labels, 1, batch), c = view(cls_loss_weights, 1, batch),
             ga = view(gt_assignment, 1, batch), d = view(pc_labels, 1, num_positive),
             e = view(pc_probs, 1, num_positive), pc = view(pc_count, 1, num_positive),
             f = view(img_cls_loss_weights, 1, num_positive), g = view(im_labels, 1, channels),
             top = at::ones({1}, at::kFloat), o = view(bottom_grad, batch, channels);
  wsl::pcl_loss_backward_cpu(a, b, c, ga, d, e, pc, f, g, top, o);
}

}  // extern "C"
