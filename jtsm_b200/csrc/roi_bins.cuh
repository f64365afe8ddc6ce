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
  // range-max-table kernel (roi_pool_rmq.cuh), channel independent
  uint32_t* bdesc;   // [R * PH * PW]  one word per output bin
  float* rscale;     // [R]            row scale (1 when the caller passes none)
  int* slow_count;   // [1]            number of bins without a fast descriptor
  int* slow_list;    // [R * PH * PW]  their ids (r * PH*PW + bin), in no particular order; bit 30 = does not fit a band
  unsigned char* bband;  // [R * PH * PW]  row band of the bin (planes larger than shared memory, RmqBands)
  // bin ids (r * PH*PW + bin) grouped by (image, band) for banded planes with <= kBandListBands bands: the banded main
  // loop walks its band's segment densely instead of filtering every bin of the call by its band
  int* band_cnt;   // [batch * kBandListBands]      bins per (image, band)
  int* band_off;   // [batch * kBandListBands + 1]  exclusive scan of band_cnt
  int* band_cur;   // [batch * kBandListBands]      fill cursors
  int* band_list;  // [R * PH * PW]
  int num_rois;    // R
};
constexpr int kBandListBands = 16;

// Row bands of a plane that is too large for shared memory (plane + 10 tables = 24 B per cell): band b covers rows
// [b * S, b * S + Hb) (the last one ends at H); consecutive bands overlap by Hb - S >= the tallest bin a RoI inside
// the map can have (ceil(H / PH) + 1), so every such bin lies inside the band floor(hstart / S).  nb == 1: the whole
// plane (S = Hb = H).  Descriptors, table entries and slow-path ranges are band relative.
struct RmqBands {
  int nb, S, Hb;
};
__host__ __device__ inline int band_of(int hs, const RmqBands& bd) { return min(max(hs, 0) / bd.S, bd.nb - 1); }
__host__ __device__ inline int band_rows(int b, int H, const RmqBands& bd) { return min(bd.Hb, H - b * bd.S); }

// Per-bin descriptor of the range-max-table kernel.  Table (kh, kw) sits in slot kh + 2*kw of the
// table block, so the 4 look-ups of a bin are base, base + db, base + dh, base + dh + db (bytes):
//   bits  0..17  base = 2 * (hs*W + ws + slot*hw)        (plane + tables <= 227 KB -> < 2^18)
//   bits 18..21  hl - hs   (dh = 2 * W * that)
//   bits 22..26  db = 2 * (wl - ws)
//   bits 27..30  table slot kh + 2*kw (the keyed kernel recovers the window's cell position from base and slot)
//   bit  31      fast path; 0 = empty bin or a shape the stored levels do not cover with 2 x 2
//                windows (the kernel then re-derives the ranges from hbins / wbins)
__device__ __forceinline__ bool rmq_axis(uint32_t packed, int& s, int& last, int& q) {
  s = range_lo(packed);
  const int len = range_hi(packed) - s;
  if (len <= 0) return false;
  q = min(3, 31 - __clz(len));
  last = range_hi(packed) - (1 << q);
  return len <= (2 << q);
}
__device__ __forceinline__ uint32_t rmq_bin_desc(uint32_t hb, uint32_t wb, int W, int hw) {
  int hs = 0, hl = 0, qh = 0, wst = 0, wl = 0, qw = 0;
  const bool okh = rmq_axis(hb, hs, hl, qh), okw = rmq_axis(wb, wst, wl, qw);
  if (!(okh && okw && abs(qh - qw) <= 1)) return 0u;
  return 0x80000000u | (2u * (uint32_t)(hs * W + wst + (qh + 2 * qw) * hw)) | ((uint32_t)(hl - hs) << 18) |
         ((2u * (uint32_t)(wl - wst)) << 22) | ((uint32_t)(qh + 2 * qw) << 27);
}

inline size_t pool_workspace_bytes(int batch, int num_rois, int pooled_h, int pooled_w, int sets) {
  size_t b = 0;
  b += align_up((size_t)sets * num_rois * pooled_h * 4, 256);
  b += align_up((size_t)sets * num_rois * pooled_w * 4, 256);
  b += align_up((size_t)num_rois * 4, 256);
  b += align_up((size_t)batch * 4 * 2 + 16, 256);
  b += 2 * align_up((size_t)num_rois * pooled_h * pooled_w * 4, 256);
  b += align_up((size_t)num_rois * 4, 256);
  b += align_up((size_t)num_rois * pooled_h * pooled_w, 256);  // bband
  b += 3 * align_up(((size_t)batch * kBandListBands + 1) * 4, 256);
  b += align_up((size_t)num_rois * pooled_h * pooled_w * 4, 256);
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
  w.slow_count = w.ungrouped + 1;
  p += align_up((size_t)batch * 4 * 2 + 16, 256);
  w.bdesc = (uint32_t*)p; p += align_up((size_t)num_rois * pooled_h * pooled_w * 4, 256);
  w.rscale = (float*)p; p += align_up((size_t)num_rois * 4, 256);
  w.slow_list = (int*)p; p += align_up((size_t)num_rois * pooled_h * pooled_w * 4, 256);
  w.bband = (unsigned char*)p; p += align_up((size_t)num_rois * pooled_h * pooled_w, 256);
  w.band_cnt = (int*)p; p += align_up(((size_t)batch * kBandListBands + 1) * 4, 256);
  w.band_off = (int*)p; p += align_up(((size_t)batch * kBandListBands + 1) * 4, 256);
  w.band_cur = (int*)p; p += align_up(((size_t)batch * kBandListBands + 1) * 4, 256);
  w.band_list = (int*)p;
  w.num_rois = num_rois;
  return w;
}

}  // namespace jtsm
