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
// One persistent CTA per SM walks a contiguous range of (plane, RoI-chunk) items, rebuilding the
// tables only when the plane changes (~3.5 builds per CTA at config 2); one thread per bin, lanes
// along consecutive bins of a RoI.  The variants tried and dropped (thread per output row with
// staged write-out, 16-byte per-bin descriptors, two bins per thread) and what each measured are
// in profiles/r1f_roi_pool_fwd_experiments.md.
#pragma once
#include "roi_bins.cuh"

namespace jtsm {

constexpr int kRmqThreads = 1024;
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

// One doubling step: dst[p] = better of src[p] and src[p + off].  No bounds logic: a partner past
// the end of the plane is clamped to the last cell, and an entry whose window would leave its row
// or the plane is simply never queried -- every query window lies inside its bin (and every valid
// entry is built from two valid halves), so only entries of in-plane windows have to be right.
__device__ __forceinline__ void rmq_build_level(const float* plane, const unsigned short* src,
                                                unsigned short* dst, int hw, int off) {
  const int last = hw - 1;
#pragma unroll 4
  for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
    const unsigned a = src[p];
    const unsigned b = src[min(p + off, last)];
    const float va = plane[a], vb = plane[b];
    const bool take_b = (vb > va) || (vb == va && b < a);
    dst[p] = (unsigned short)(take_b ? b : a);
  }
}

// plane (fp32) + the 10 tables of one (image, channel) plane; all threads of the CTA call it.
template <typename T>
__device__ __forceinline__ void rmq_build_plane(const T* __restrict__ src, float* plane, unsigned short* tabs,
                                                int hw, int W) {
#define TAB(s) (tabs + (size_t)(s) * hw)
  for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
    plane[p] = Elem<T>::to_f(src[p]);
    TAB(0)[p] = (unsigned short)p;  // identity level (0,0)
  }
  __syncthreads();
  rmq_build_level(plane, TAB(0), TAB(2), hw, 1);      // (0,1)
  rmq_build_level(plane, TAB(0), TAB(1), hw, W);      // (1,0)
  __syncthreads();
  rmq_build_level(plane, TAB(2), TAB(3), hw, W);      // (1,1)
  __syncthreads();
  rmq_build_level(plane, TAB(3), TAB(5), hw, 2);      // (1,2)
  rmq_build_level(plane, TAB(3), TAB(4), hw, 2 * W);  // (2,1)
  __syncthreads();
  rmq_build_level(plane, TAB(5), TAB(6), hw, 2 * W);  // (2,2)
  __syncthreads();
  rmq_build_level(plane, TAB(6), TAB(8), hw, 4);      // (2,3)
  rmq_build_level(plane, TAB(6), TAB(7), hw, 4 * W);  // (3,2)
  __syncthreads();
  rmq_build_level(plane, TAB(8), TAB(9), hw, 4 * W);  // (3,3)
  __syncthreads();
#undef TAB
}

// Slow path of one bin (empty, elongated or > 16-cell ranges) from the packed [start, end) ranges:
// nearest stored level, as many windows as needed.  Empty bin -> value 0, index 0xFFFFFFFF.
__device__ __noinline__ RmqCand rmq_query_slow(const float* plane, const unsigned short* tabs, int hw, int W,
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

// 2 x 2 windows of a stored level: byte offsets base, +db, +dh, +dh+db into the table block.
__device__ __forceinline__ RmqCand rmq_query_fast(const float* plane, const unsigned short* tabs, unsigned base,
                                                  unsigned dh, unsigned db) {
  const char* tb = (const char*)tabs + base;
  const unsigned ia = *(const unsigned short*)(tb);
  const unsigned ib = *(const unsigned short*)(tb + db);
  const unsigned ic = *(const unsigned short*)(tb + dh);
  const unsigned id = *(const unsigned short*)(tb + dh + db);
  return rmq_best(rmq_best(rmq_load(plane, ia), rmq_load(plane, ib)),
                  rmq_best(rmq_load(plane, ic), rmq_load(plane, id)));
}

struct RmqItem {
  int n, c, r0, nr;
};
// (plane, chunk) item -> image, channel and RoI range; nr = 0 when the image has no RoIs
__device__ __forceinline__ RmqItem rmq_item(int item, int chunks, int channels, const PoolWorkspace& ws, int& nc) {
  RmqItem it;
  nc = item / chunks;
  const int chunk = item - nc * chunks;
  it.n = nc / channels;
  it.c = nc - it.n * channels;
  const int first = ws.first[it.n];
  const int last = -ws.neg_last[it.n];
  it.r0 = 0;
  it.nr = 0;
  if (first <= last) {
    const int len = last - first + 1;
    it.r0 = first + (int)(((long long)len * chunk) / chunks);
    it.nr = first + (int)(((long long)len * (chunk + 1)) / chunks) - it.r0;
  }
  return it;
}

template <typename T, typename A>
__global__ void __launch_bounds__(kRmqThreads, 1)
roi_pool_fwd_rmq_kernel(const T* __restrict__ input, int channels, int H, int W, int PH, int PW,
                        int chunks, int num_items, PoolWorkspace ws, T* __restrict__ output,
                        A* __restrict__ argmax) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int hw = H * W;
  float* plane = (float*)smem_raw;
  unsigned short* tabs = (unsigned short*)(plane + hw);
  const int PP = PH * PW, cpp = channels * PP;
  const unsigned w2 = 2u * (unsigned)W;
  const bool ungrouped = *ws.ungrouped != 0;
  // (RoI, bin) of this thread's output and the step of one round (kRmqThreads outputs); the output
  // offset rl * cpp + bin is stepped along with them
  const int step_r = kRmqThreads / PP, step_b = kRmqThreads - step_r * PP;
  const int step_o = step_r * cpp + step_b, wrap_o = cpp - PP;
  const int rl_init = threadIdx.x / PP, bin_init = threadIdx.x - rl_init * PP;

  // contiguous item range of this CTA
  const int per = num_items / gridDim.x, rem = num_items % gridDim.x;
  const int i0 = blockIdx.x * per + min((int)blockIdx.x, rem);
  const int i1 = i0 + per + ((int)blockIdx.x < rem ? 1 : 0);
  int cur_plane = -1;

  for (int item = i0; item < i1; ++item) {
    int nc;
    const RmqItem it = rmq_item(item, chunks, channels, ws, nc);
    if (nc != cur_plane) {
      __syncthreads();  // previous item's queries are done with the old tables
      cur_plane = nc;
      rmq_build_plane(input + (size_t)nc * hw, plane, tabs, hw, W);
    }
    // All indices are 32-bit offsets from the kernel's own pointer arguments (the launcher checks the
    // sizes), so the loop carries five integers and no 64-bit pointers.
    unsigned i = (unsigned)(it.r0 * PP) + threadIdx.x;           // bin id
    const unsigned i_end = (unsigned)((it.r0 + it.nr) * PP);
    unsigned r = (unsigned)(it.r0 + rl_init);                    // RoI of bin i
    int bin = bin_init;                                          // i - r * PP
    unsigned o = (r * (unsigned)channels + (unsigned)it.c) * (unsigned)PP + (unsigned)bin;  // output element
    // descriptor and row scale are loaded one round ahead (they come from L2: little L1 is left next to
    // 180+ KB of shared memory), so their latency hides behind the look-ups of the current round
    unsigned xn = 0u;
    float scn = 0.f;
    if (i < i_end) xn = __ldg(ws.bdesc + i), scn = __ldg(ws.rscale + r);
    while (i < i_end) {
      const unsigned x = xn;  // x = 0: slow bin, left to the list below (the look-ups then run on offset 0)
      const float sc = scn;
      bool act = (int)x < 0;
      if (ungrouped && act) act = __ldg(ws.batch_idx + r) == it.n;
      const unsigned o_cur = o;
      i += kRmqThreads;
      bin += step_b;
      r += step_r;
      o += step_o;
      if (bin >= PP) bin -= PP, ++r, o += wrap_o;
      if (i < i_end) xn = __ldg(ws.bdesc + i), scn = __ldg(ws.rscale + r);
      const RmqCand best = rmq_query_fast(plane, tabs, x & 0x3FFFFu, ((x >> 18) & 15u) * w2, (x >> 22) & 31u);
      const float v = __fmul_rn(best.v, sc);
      if (act) {
        output[o_cur] = Elem<T>::from_f(v);
        argmax[o_cur] = Arg<A>::enc((int)best.i);
      }
    }
    // the few bins without a fast descriptor (empty, elongated, > 16 cells: ~2%), listed by the prep
    // kernel so that no warp of the main loop diverges into the long path; the list is short, every
    // item filters it for its own RoI range
    const int ns = __ldg(ws.slow_count);
    for (int j = threadIdx.x; j < ns; j += kRmqThreads) {
      const int id = __ldg(ws.slow_list + j);
      const int rr = id / PP, b = id - rr * PP;
      if (rr < it.r0 || rr >= it.r0 + it.nr) continue;
      if (ungrouped && __ldg(ws.batch_idx + rr) != it.n) continue;
      const int ph = b / PW, pw = b - ph * PW;
      const RmqCand best = rmq_query_slow(plane, tabs, hw, W, __ldg(ws.hbins + (size_t)rr * PH + ph),
                                          __ldg(ws.wbins + (size_t)rr * PW + pw));
      const unsigned so = ((unsigned)rr * (unsigned)channels + (unsigned)it.c) * (unsigned)PP + (unsigned)b;
      output[so] = Elem<T>::from_f(__fmul_rn(best.v, __ldg(ws.rscale + rr)));
      argmax[so] = Arg<A>::enc((int)best.i);
    }
  }
}

}  // namespace jtsm
