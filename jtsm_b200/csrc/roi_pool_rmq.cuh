// RoIPool forward with per-plane 2-D range-max index tables ("rmq" kernels).
//
// Why: the direct per-bin scan reads RoI-area x ~1.6 cells per (RoI, channel) -- 1.07 G shared
// memory reads per step at config 2 -- and is LSU/issue bound at ~880 us (profiles/r1a), 20x off
// the HBM time of the 200 MB output.  A max over a rectangle is an idempotent range query, so a
// sparse table answers it with a constant number of look-ups:
//
//   T[kh][kw][h][w] = row-major index of the FIRST maximum of plane[h .. h+2^kh) x [w .. w+2^kw)
//
// (uint16 absolute index; "first in row-major order" is exactly the reference's strict `>` scan
// order, ROILoopPool_cpu.cpp:55-66, so argmax stays bit-exact).  A bin [hs,he) x [ws,we) is
// covered by 2 x 2 overlapping windows of the level floor(log2(len)); the answer is the best of
// the 4 candidates (value desc, then index asc).  Only levels with |kh - kw| <= 1 are stored
// (10 tables incl. identity, 24 B per cell in total with the fp32 plane -> 180 KB at 75x100);
// more elongated (or empty, or > 16-cell) bins -- ~2% -- go through a list the prep kernel makes
// and a slow path after the main loop, so no warp of the main loop diverges.
//
// Everything channel independent about a bin -- the byte offset of its first look-up, the two
// window strides and whether it is a fast bin -- is one 32-bit descriptor written by the prep
// kernel (roi_bins.cuh) and read with one coalesced load, one round ahead; the look-up chain of
// the main loop is branch free and indexes only with 32-bit offsets from the kernel arguments.
//
// One persistent CTA per SM walks (plane, RoI-chunk) items, rebuilding the tables only when the plane
// changes (~3.5 builds per CTA at config 2); one thread per bin, lanes along consecutive bins of a RoI.
// The kernel frame (schedule, bulk-copy plane fetch, unpack, main loop, slow pass) is in
// roi_pool_rmqk.cuh, shared with the keyed table format defined there; this file holds the index-table
// format (build + queries) and what both formats share.  The variants tried and dropped (thread per
// output row with staged write-out, 16-byte per-bin descriptors, keyed entries) and what each measured
// are in profiles/r1f_roi_pool_fwd_experiments.md and profiles/r2_roi_pool_fwd_keyed.md.
#pragma once
#include "roi_bins.cuh"

namespace jtsm {

constexpr int kRmqThreads = 1024;
constexpr int kRmqBinsPerThread = 2;  // independent look-up chains per thread and round
constexpr int kRmqTables = 10;   // slot = kh + 2*kw for |kh-kw| <= 1, kh,kw in 0..3; slot 0 = identity
constexpr int kRmqMaxLevel = 3;  // windows of 1,2,4,8 cells per axis

__host__ __device__ inline size_t rmq_smem_bytes(int hw) {
  return (size_t)hw * 4 + (size_t)kRmqTables * hw * 2 + 16;
}

__device__ __forceinline__ int rmq_slot(int kh, int kw) { return kh + 2 * kw; }

// exact p / d for p < 2^32 / d (magic = ceil(2^32 / d))
__device__ __forceinline__ int fast_div(int p, unsigned magic) { return (int)__umulhi((unsigned)p, magic); }

// (value, index) candidate; better = larger value, then smaller index (= first in row-major order)
struct RmqCand {
  float v;
  unsigned i;
};
__device__ __forceinline__ RmqCand rmq_load(const float* plane, unsigned idx) { return {plane[idx], idx}; }
__device__ __forceinline__ RmqCand rmq_best(const RmqCand a, const RmqCand b) {
  const bool take_b = (b.v > a.v) || (b.v == a.v && b.i < a.i);
  return take_b ? b : a;
}

// MOIPool mode of the range-max kernel (wsl_pool.cu): a bin whose cells are ALL switched on in the superpixel mask is a
// plain RoIPool bin and keeps its fast descriptor; this kernel answers exactly those (main loop only, no slow pass)
// and the visit-list kernel (moi_pool_fwd_list_kernel, two planes per CTA) answers the rest.  It only runs when at
// most `gate_pct` percent of the bins are left to the lists (decided on the device, no host sync; the list kernel
// makes the same test and then answers every bin itself).
struct MoiAux {
  const float* mask_scale;  // [R]  P*P / (#non-empty bins + 1); the row scale stays in ws.rscale
  const int* slow_total;    // [1] number of bins left to the visit lists (the gate)
  int total_bins;
  int gate_pct;
  int zero_floor;           // ROILoopPool: values <= 0 give (0, -1) -- its running maximum starts at 0
};
__device__ __forceinline__ bool moi_rmq_gate(const int* slow_count, int total_bins, int gate_pct) {
  return (long long)__ldg(slow_count) * 100 <= (long long)total_bins * gate_pct;
}

// One build phase: from level (k, k) in `src` (window 2^k x 2^k, s = 2^k) to the three levels (k, k+1), (k+1, k)
// and (k+1, k+1): the four (k, k) windows at p, p + s, p + s*W, p + s*W + s give all three with 4 index loads,
// 4 value loads and 3 comparisons.  kIdentity: level (0, 0) is the cell itself, no index loads.
// No bounds logic: a partner past the end of the plane is clamped to the last cell, and an entry whose window
// would leave its row or the plane is simply never queried -- every query window lies inside its bin (and every
// valid entry is built from valid ones), so only entries of in-plane windows have to be right.
template <bool kIdentity>
__device__ __forceinline__ void rmq_build_phase(const float* plane, const unsigned short* src, unsigned short* d_w,
                                                unsigned short* d_h, unsigned short* d_hw, int hw, int W, int s) {
  const int last = hw - 1;
#pragma unroll 2
  for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
    const int pb = min(p + s, last), pc = min(p + s * W, last), pd = min(p + s * W + s, last);
    const RmqCand a = rmq_load(plane, kIdentity ? (unsigned)p : (unsigned)src[p]);
    const RmqCand b = rmq_load(plane, kIdentity ? (unsigned)pb : (unsigned)src[pb]);
    const RmqCand c = rmq_load(plane, kIdentity ? (unsigned)pc : (unsigned)src[pc]);
    const RmqCand d = rmq_load(plane, kIdentity ? (unsigned)pd : (unsigned)src[pd]);
    const RmqCand ab = rmq_best(a, b), cd = rmq_best(c, d);
    d_w[p] = (unsigned short)ab.i;                 // (k, k+1): wider
    d_h[p] = (unsigned short)rmq_best(a, c).i;     // (k+1, k): taller
    d_hw[p] = (unsigned short)rmq_best(ab, cd).i;  // (k+1, k+1)
  }
}

// The 9 non-identity tables of the plane in `plane` (slot 0, the identity level, is written once per kernel);
// all threads of the CTA call it, after a barrier that makes the plane visible.  Slot of level (kh, kw) = kh + 2*kw.
__device__ __forceinline__ void rmq_build_tables(const float* plane, unsigned short* tabs, int hw, int W) {
#define TAB(s) (tabs + (size_t)(s) * hw)
  rmq_build_phase<true>(plane, TAB(0), TAB(2), TAB(1), TAB(3), hw, W, 1);   // (0,1) (1,0) (1,1)
  __syncthreads();
  rmq_build_phase<false>(plane, TAB(3), TAB(5), TAB(4), TAB(6), hw, W, 2);  // (1,2) (2,1) (2,2)
  __syncthreads();
  rmq_build_phase<false>(plane, TAB(6), TAB(8), TAB(7), TAB(9), hw, W, 4);  // (2,3) (3,2) (3,3)
  __syncthreads();
#undef TAB
}

// Slow path of one bin (empty, elongated or > 16-cell ranges) from the packed [start, end) ranges:
// nearest stored level, as many windows as needed.  Empty bin -> value 0, index 0xFFFFFFFF.
static __device__ __noinline__ RmqCand rmq_query_slow(const float* plane, const unsigned short* tabs, int hw, int W,
                                               uint32_t hb, uint32_t wb) {
  const int hs = range_lo(hb), he = range_hi(hb), wst = range_lo(wb), wen = range_hi(wb);
  const int lh = he - hs, lw = wen - wst;
  if (lh <= 0 || lw <= 0) return {0.f, 0xFFFFFFFFu};
  const int qh = min(kRmqMaxLevel, 31 - __clz(lh));
  const int qw = min(kRmqMaxLevel, 31 - __clz(lw));
  const int kh = min(qh, qw + 1), kw = min(qw, qh + 1);
  const unsigned short* tab = tabs + (size_t)rmq_slot(kh, kw) * hw;
  const int sh = 1 << kh, sw = 1 << kw;
  const int hl = he - sh, wl = wen - sw;
  RmqCand best = rmq_load(plane, tab[hs * W + wst]);
  for (int hh = hs;; hh += sh) {
    const int h2 = min(hh, hl);
    for (int ww = wst;; ww += sw) {
      const int w2 = min(ww, wl);
      best = rmq_best(best, rmq_load(plane, tab[h2 * W + w2]));
      if (ww >= wl) break;
    }
    if (hh >= hl) break;
  }
  return best;
}

// Shared-memory loads by 32-bit shared-window address: the address arithmetic of the hot loop stays one add
// per look-up (the compiler otherwise re-derives the window base from the CTA rank every round).
__device__ __forceinline__ unsigned lds_u16(unsigned addr) {
  unsigned short v;
  asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds_f32(unsigned addr) {
  float v;
  asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// 2 x 2 windows of a stored level: table look-ups at base, +db, +dh, +dh+db (shared-window byte addresses),
// then the 4 candidate values from the plane at plane_s.
__device__ __forceinline__ RmqCand rmq_query_fast(unsigned plane_s, unsigned base, unsigned dh, unsigned db) {
  const unsigned ia = lds_u16(base), ib = lds_u16(base + db), ic = lds_u16(base + dh), id = lds_u16(base + dh + db);
  const RmqCand a{lds_f32(plane_s + 4u * ia), ia}, b{lds_f32(plane_s + 4u * ib), ib};
  const RmqCand c{lds_f32(plane_s + 4u * ic), ic}, d{lds_f32(plane_s + 4u * id), id};
  return rmq_best(rmq_best(a, b), rmq_best(c, d));
}

struct RmqItem {
  int n, c, r0, nr;
};
// (plane, chunk) -> image, channel and RoI range; nr = 0 when the image has no RoIs.  32-bit arithmetic only
// (the launcher bounds num_rois * chunks).
__device__ __forceinline__ RmqItem rmq_item(int nc, int chunk, int chunks, int channels, const PoolWorkspace& ws) {
  RmqItem it;
  it.n = nc / channels;
  it.c = nc - it.n * channels;
  const int first = ws.first[it.n];
  const int last = -ws.neg_last[it.n];
  it.r0 = 0;
  it.nr = 0;
  if (first <= last) {
    const unsigned len = (unsigned)(last - first + 1);
    it.r0 = first + (int)(len * (unsigned)chunk / (unsigned)chunks);
    it.nr = first + (int)(len * (unsigned)(chunk + 1) / (unsigned)chunks) - it.r0;
  }
  return it;
}

}  // namespace jtsm
