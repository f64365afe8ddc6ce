// extern "C" driver around the reference's OWN CPU RoIAlign (detectron2/layers/csrc/ROIAlign/ROIAlign_cpu.cpp:401-506,
// declared in ROIAlign.h:7-25), compiled in place from /root/reference by oracle/Makefile.ref against this image's
// ATen.  Plain pointers in, plain pointers out; the ATen tensors are views made with at::from_blob.
// TEST INFRASTRUCTURE ONLY: pins oracle/jtsm_oracle.c's RoIAlign restatement; never linked into the product.
#include <ATen/ATen.h>

#include <cstring>

#include "ROIAlign/ROIAlign.h"

extern "C" {

void ref_roi_align_forward(const float* input, int n, int c, int h, int w, const float* rois, int r,
                           float spatial_scale, int ph, int pw, int sampling_ratio, int aligned, float* out) {
  at::Tensor x = at::from_blob(const_cast<float*>(input), {n, c, h, w}, at::kFloat);
  at::Tensor b = at::from_blob(const_cast<float*>(rois), {r, 5}, at::kFloat);
  at::Tensor y = detectron2::ROIAlign_forward_cpu(x, b, spatial_scale, ph, pw, sampling_ratio, aligned != 0);
  y = y.contiguous();
  std::memcpy(out, y.data_ptr<float>(), (size_t)y.numel() * sizeof(float));
}

void ref_roi_align_backward(const float* grad, const float* rois, int r, float spatial_scale, int ph, int pw, int n,
                            int c, int h, int w, int sampling_ratio, int aligned, float* grad_input) {
  at::Tensor g = at::from_blob(const_cast<float*>(grad), {r, c, ph, pw}, at::kFloat);
  at::Tensor b = at::from_blob(const_cast<float*>(rois), {r, 5}, at::kFloat);
  at::Tensor gi =
      detectron2::ROIAlign_backward_cpu(g, b, spatial_scale, ph, pw, n, c, h, w, sampling_ratio, aligned != 0);
  gi = gi.contiguous();
  std::memcpy(grad_input, gi.data_ptr<float>(), (size_t)gi.numel() * sizeof(float));
}

}  // extern "C"
