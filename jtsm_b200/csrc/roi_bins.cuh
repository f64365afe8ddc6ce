// RoI -> integer bin-edge tables shared by the max-pool family
// (RoIPool, ROILoopPool, MOIPool).  The tables are channel independent, so they
// are computed once per call (R * P entries) instead of once per output element
// as the reference kernels do (ROILoopPool_cuda.cu:77-104, MOIPool_cuda.cu:240-260).
#pragma once
#include "common.cuh"

namespace jtsm {

// Packed [start, end) cell range: low 16 bits = start, high 16 bits = end.
__device__ __forceinline__ uint32_t pack_range(int s, int e) {
  return (uint32_t)(s & 0xFFFF) | ((uint32_t)(e & 0xFFFF) << 16);
}
__device__ __forceinline__ int range_lo(uint32_t p) { return (int)(p & 0xFFFFu); }
__device__ __forceinline__ int range_hi(uint32_t p) { return (int)(p >> 16); }

// Integer RoI rectangle in feature-map cells, exactly as the reference:
//   roi_start = round(x * spatial_scale)      (roundf: half away from zero)
// ROILoopPool_cpu.cpp:29-32.
struct RoiRect {
  int sw, sh, ew, eh;
};
__device__ __forceinline__ RoiRect roi_rect(float x1, float y1, float x2, float y2, float scale) {
  RoiRect r;
  r.sw = (int)roundf(__fmul_rn(x1, scale));
  r.sh = (int)roundf(__fmul_rn(y1, scale));
  r.ew = (int)roundf(__fmul_rn(x2, scale));
  r.eh = (int)roundf(__fmul_rn(y2, scale));
  return r;
}

// Bin p of an axis of `roi_len = max(end - start + 1, 1)` cells split in `P`:
//   start = floor(p * bin), end = ceil((p + 1) * bin), bin = roi_len / P (float),
// then offset by roi_start and clipped to [0, limit]   (ROILoopPool_cpu.cpp:35-51).
__device__ __forceinline__ uint32_t bin_range(int p, int roi_start, int roi_len, int P, int limit) {
  const float bin = __fdiv_rn((float)roi_len, (float)P);
  int s = (int)floorf(__fmul_rn((float)p, bin));
  int e = (int)ceilf(__fmul_rn((float)(p + 1), bin));
  s = min(max(s + roi_start, 0), limit);
  e = min(max(e + roi_start, 0), limit);
  return pack_range(s, e);
}

// Workspace carved up for the max-pool family.
struct PoolWorkspace {
  uint32_t* hbins;   // [sets * R * P]
  uint32_t* wbins;   // [sets * R * P]
  int* batch_idx;    // [R]
  int* first;        // [batch]   min roi index of image n   (INT-ish max if none)
  int* neg_last;     // [batch]   -(max roi index of image n)
  int* ungrouped;    // [1]       != 0 when batch indices are not non-decreasing in r
  // range-max-table kernel descriptors (roi_pool_rmq.cuh); byte offsets into the table block
  uint4* hdesc;      // [R * P]  {2*(hs*W + qh*hw), 2*(hl*W + qh*hw), row_scale bits, hgood << 28 | batch & 0xFFFFFF}
  uint32_t* wdesc;   // [R * wstride]  2*(ws + 2*qw*hw) | 2*(wl - ws) << 20 | wgood << 28 ; wstride = rmq_wstride(P)
};

// descriptor of one axis range [s, e): level q = floor(log2(len)) capped at 3, start of the last
// window l = e - 2^q.  x/y = byte offsets of the two windows inside table slot q*slot_mult
// (cell stride `mult`); `good` = 4-bit mask, 0 when the range is empty or two windows of the level
// do not cover it: for the h axis the set of w levels stored next to q (|qh - qw| <= 1), for the
// w axis the single bit 1 << q -- a bin takes the fast path iff hgood & wgood != 0.
__device__ __forceinline__ void rmq_axis_desc(uint32_t packed, int mult, int slot_mult, int hw, bool h_axis,
                                              uint32_t& x, uint32_t& y, uint32_t& good) {
  const int s = range_lo(packed), e = range_hi(packed);
  const int len = e - s;
  x = y = good = 0;
  if (len <= 0) return;
  const int q = min(3, 31 - __clz(len));
  if (len > (2 << q)) return;
  const int base = q * slot_mult * hw;
  x = 2u * (uint32_t)(s * mult + base);
  y = 2u * (uint32_t)((e - (1 << q)) * mult + base);
  good = h_axis ? ((0x7u << q) >> 1) & 0xFu : 1u << q;
}

// words per RoI in wdesc: PW rounded up to a multiple of 4 so a row loads as 128-bit words
__host__ __device__ inline int rmq_wstride(int pooled_w) { return (pooled_w + 3) & ~3; }

inline size_t pool_workspace_bytes(int batch, int num_rois, int pooled_h, int pooled_w, int sets) {
  size_t b = 0;
  b += align_up((size_t)sets * num_rois * pooled_h * 4, 256);
  b += align_up((size_t)sets * num_rois * pooled_w * 4, 256);
  b += align_up((size_t)num_rois * 4, 256);
  b += align_up((size_t)batch * 4 * 2 + 16, 256);
  b += align_up((size_t)num_rois * pooled_h * 16, 256);
  b += align_up((size_t)num_rois * rmq_wstride(pooled_w) * 4, 256);
  return b + 256;
}

inline PoolWorkspace carve_pool_workspace(void* ws, int batch, int num_rois, int pooled_h, int pooled_w, int sets) {
  char* p = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  PoolWorkspace w;
  w.hbins = (uint32_t*)p; p += align_up((size_t)sets * num_rois * pooled_h * 4, 256);
  w.wbins = (uint32_t*)p; p += align_up((size_t)sets * num_rois * pooled_w * 4, 256);
  w.batch_idx = (int*)p; p += align_up((size_t)num_rois * 4, 256);
  w.first = (int*)p;
  w.neg_last = w.first + batch;
  w.ungrouped = w.neg_last + batch;
  p += align_up((size_t)batch * 4 * 2 + 16, 256);
  w.hdesc = (uint4*)p; p += align_up((size_t)num_rois * pooled_h * 16, 256);
  w.wdesc = (uint32_t*)p;
  return w;
}

}  // namespace jtsm
