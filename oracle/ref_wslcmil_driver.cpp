// extern "C" driver around the reference's OWN CPU ROIMerge / ROILabel (projects/WSL/wsl/layers/csrc/ROIMerge/
// ROIMerge_cpu.cpp:33-232 forward, :234-287 backward; ROILabel/ROILabel_cpu.cpp:16-194), compiled in place by
// oracle/Makefile.ref against this image's ATen.  TEST INFRASTRUCTURE ONLY: pins oracle/ref_cmil.py's restatements; never
// linked into the product.
//
// ROILabel visits the RoIs in the order of std::random_shuffle seeded with std::time(0) (ROILabel_cpu.cpp:108-117).  This
// library is linked -Bsymbolic and defines time() itself, so the reference's srand(time(0)) always seeds with kSeed and
// ref_roi_label_perm() can hand the test the very permutation the reference's call will draw (same libstdc++
// random_shuffle over the same rand() stream).
#include <ATen/ATen.h>

#include <algorithm>
#include <cstdlib>
#include <ctime>
#include <vector>

#include "ROIMerge/ROIMerge.h"
#include "ROILabel/ROILabel.h"

static const unsigned kSeed = 20240229u;
extern "C" time_t time(time_t* t) {
  if (t) *t = (time_t)kSeed;
  return (time_t)kSeed;
}

static at::Tensor fview(const float* p, long rows, long cols) {
  return at::from_blob(const_cast<float*>(p), {rows, cols}, at::kFloat);
}

extern "C" {

// P int32[8] in/out (debug_info, display, cur_iter, max_epoch, size_epoch, acc_num_top_id, acc_max_clique, acc_min_clique).
// MC / MD receive num_id rows (caller provides num_roi rows); returns num_id.
int ref_roi_merge_forward(const float* S, const float* J, const float* C, const float* D, int* P, int num_roi,
                          int num_class, float* MC, float* MD, int* I, int* IC) {
  at::Tensor p = at::from_blob(P, {8}, at::kInt);
  auto out = wsl::ROIMerge_forward_cpu(fview(S, num_roi, 1), fview(J, num_roi, num_roi), fview(C, num_roi, num_class),
                                       fview(D, num_roi, num_class), p);
  at::Tensor mc = std::get<0>(out).contiguous(), md = std::get<1>(out).contiguous();
  at::Tensor i = std::get<2>(out).contiguous(), ic = std::get<3>(out).contiguous();
  const long num_id = mc.size(0);
  std::copy(mc.data_ptr<float>(), mc.data_ptr<float>() + num_id * num_class, MC);
  std::copy(md.data_ptr<float>(), md.data_ptr<float>() + num_id * num_class, MD);
  std::copy(i.data_ptr<int>(), i.data_ptr<int>() + num_roi, I);
  std::copy(ic.data_ptr<int>(), ic.data_ptr<int>() + num_roi, IC);
  return (int)num_id;
}

void ref_roi_merge_backward(const float* C, const float* D, const float* GMC, const float* GMD, const int* I, const int* IC,
                            int num_roi, int num_class, int num_id, float* GC, float* GD) {
  at::Tensor i = at::from_blob(const_cast<int*>(I), {num_roi}, at::kInt);
  at::Tensor ic = at::from_blob(const_cast<int*>(IC), {num_roi}, at::kInt);
  auto out = wsl::ROIMerge_backward_cpu(fview(C, num_roi, num_class), fview(D, num_roi, num_class),
                                        fview(GMC, num_id, num_class), fview(GMD, num_id, num_class), i, ic);
  at::Tensor gc = std::get<0>(out).contiguous(), gd = std::get<1>(out).contiguous();
  std::copy(gc.data_ptr<float>(), gc.data_ptr<float>() + (long)num_roi * num_class, GC);
  std::copy(gd.data_ptr<float>(), gd.data_ptr<float>() + (long)num_roi * num_class, GD);
}

// The permutation ROILabel_forward_cpu will visit the RoIs in (see the header comment).
void ref_roi_label_perm(int num_roi, int* perm) {
  std::srand(unsigned(std::time(0)));
  std::vector<int> v;
  for (int n = 0; n < num_roi; n++) v.push_back(n);
  std::random_shuffle(v.begin(), v.end());
  std::copy(v.begin(), v.end(), perm);
}

// P float[14] in/out (ROILabel_cpu.cpp:44-58).  CW may be NULL.
void ref_roi_label_forward(const float* S, const float* U, const float* L, const float* CW, float* P, int num_roi,
                           int num_class_s, int num_class, int* RL, float* RW) {
  at::Tensor p = at::from_blob(P, {14}, at::kFloat);
  at::Tensor cw = CW ? fview(CW, 1, num_class) : at::Tensor();
  if (!CW) cw = at::empty({0}, at::kFloat);
  auto out = wsl::ROILabel_forward_cpu(fview(S, num_roi, num_class_s), fview(U, num_roi, num_roi), fview(L, 1, num_class),
                                       cw, p);
  at::Tensor rl = std::get<0>(out).contiguous(), rw = std::get<1>(out).contiguous();
  std::copy(rl.data_ptr<int>(), rl.data_ptr<int>() + num_roi, RL);
  std::copy(rw.data_ptr<float>(), rw.data_ptr<float>() + num_roi, RW);
}

}  // extern "C"
