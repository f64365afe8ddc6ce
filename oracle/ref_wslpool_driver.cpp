// extern "C" driver around the reference's OWN ROILoopPool / MOIPool kernels, compiled as serial host code
// (oracle/Makefile.ref).  It restates only what the ATen launchers do around the kernels:
//   ROILoopPool_forward_cuda  projects/WSL/wsl/layers/csrc/ROILoopPool/ROILoopPool_cuda.cu:252-313
//       output / argmax = zeros [3R, C, PH, PW]; one kernel over R*C*PH*PW elements; context ratio 1.8 (:309)
//   MOIPool_forward_cuda      projects/WSL/wsl/layers/csrc/MOIPool/MOIPool_cuda.cu:342-433
//       output / argmax = zeros; mois = zeros int [R, H, W]; MoIForward over R*H*W, then MoIPoolForward2
//   *_backward_cuda           ROILoopPool_cuda.cu:315-386, MOIPool_cuda.cu:435-506 (grad_input = zeros, atomicAdd)
// TEST INFRASTRUCTURE ONLY: used by tests/ to pin oracle/jtsm_oracle.c; never linked into the product.
#include <cstring>
#include <vector>

#include "ref_host_cuda.h"
// kernel templates of the two reference files (everything before `namespace wsl`), produced by Makefile.ref
#include "ref_roi_loop_pool_kernels.inc"
#include "ref_moi_pool_kernels.inc"

extern "C" {

void ref_roi_loop_pool_forward(const float* input, const float* rois, int num_rois, int channels, int height,
                               int width, int ph, int pw, float spatial_scale, float* output, int* argmax) {
  size_t n3 = (size_t)3 * num_rois * channels * ph * pw;
  std::memset(output, 0, n3 * sizeof(float));
  std::memset(argmax, 0, n3 * sizeof(int));
  RoILoopPoolForward<float>(num_rois * channels * ph * pw, input, spatial_scale, channels, height, width, ph, pw,
                            rois, output, argmax, num_rois, 1.8);
}

void ref_roi_loop_pool_backward(const float* grad_output, const int* argmax, const float* rois, int num_rois,
                                int batch, int channels, int height, int width, int ph, int pw,
                                float spatial_scale, float* grad_input) {
  std::memset(grad_input, 0, (size_t)batch * channels * height * width * sizeof(float));
  // contiguous grad [3R, C, PH, PW]: strides as the launcher reads them (ROILoopPool_cuda.cu:358-361)
  RoILoopPoolBackward<float>(3 * num_rois * channels * ph * pw, grad_output, argmax, num_rois, spatial_scale,
                             channels, height, width, ph, pw, grad_input, rois, channels * ph * pw, ph * pw, pw, 1);
}

void ref_moi_pool_forward(const float* input, const float* rois, int num_rois, const int* oh_labels, int len_labels,
                          const int* superpixels, int height_sp, int width_sp, int channels, int height, int width,
                          int ph, int pw, float spatial_scale, float* output, int* argmax, int* mois_out) {
  size_t n = (size_t)num_rois * channels * ph * pw;
  std::memset(output, 0, n * sizeof(float));
  std::memset(argmax, 0, n * sizeof(int));
  std::vector<int> mois((size_t)num_rois * height * width, 0);
  MoIForward<float>(num_rois * height * width, spatial_scale, height, width, height_sp, width_sp, rois, mois.data(),
                    oh_labels, superpixels, len_labels);
  MoIPoolForward2<float>((int)n, input, spatial_scale, channels, height, width, ph, pw, rois, output, argmax,
                         mois.data());
  if (mois_out) std::memcpy(mois_out, mois.data(), mois.size() * sizeof(int));
}

void ref_moi_pool_backward(const float* grad_output, const int* argmax, const float* rois, int num_rois, int batch,
                           int channels, int height, int width, int ph, int pw, float spatial_scale,
                           float* grad_input) {
  std::memset(grad_input, 0, (size_t)batch * channels * height * width * sizeof(float));
  RoIPoolBackward<float>(num_rois * channels * ph * pw, grad_output, argmax, num_rois, spatial_scale, channels,
                         height, width, ph, pw, grad_input, rois, channels * ph * pw, ph * pw, pw, 1);
}

}  // extern "C"
