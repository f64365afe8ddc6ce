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
// more elongated bins use the nearest stored level and a few more windows.
//
// Table (kh, kw) lives in slot kh + 2 * kw, so the byte address of a look-up is the SUM of a
// per-(RoI, ph) part 2*(h*W + kh*hw) and a per-(RoI, pw) part 2*(w + 2*kw*hw): the prep kernel
// stores both parts (the w part packed in one word), and whether the pair of levels is a stored one is a single AND of two masks.
//
// One persistent CTA per SM walks a contiguous range of (plane, RoI-chunk) items, rebuilding the
// tables only when the plane changes (~3.5 builds per CTA at config 2); one thread per bin, lanes
// along consecutive bins of a RoI.  (A one-thread-per-output-row variant with staged write-out
// needed 2.5x fewer instructions but 25% more shared-memory wavefronts -- lanes on different rows
// collide on banks -- and measured 418 us vs 362 us, so it is not kept: profiles/r1f.)
#pragma once
#include "roi_bins.cuh"

namespace jtsm {

constexpr int kRmqThreads = 1024;
constexpr int kRmqTables = 10;   // slot = kh + 2*kw for |kh-kw| <= 1, kh,kw in 0..3; slot 0 = identity
constexpr int kRmqMaxLevel = 3;  // windows of 1,2,4,8 cells per axis
constexpr unsigned kRmqGoodMask = 0xF0000000u;

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

template <bool kVertical>
__device__ __forceinline__ void rmq_build_level(const float* plane, const unsigned short* src,
                                                unsigned short* dst, int hw, int W, unsigned magic_w, int half) {
  for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
    const unsigned a = src[p];
    unsigned r = a;
    bool has;
    if (kVertical) {
      has = p + half * W < hw;
    } else {
      const int w = p - fast_div(p, magic_w) * W;
      has = w + half < W;
    }
    if (has) {
      const unsigned b = src[p + (kVertical ? half * W : half)];
      if (b != a) r = rmq_best(rmq_load(plane, a), rmq_load(plane, b)).i;
    }
    dst[p] = (unsigned short)r;
  }
}

// plane (fp32) + the 10 tables of one (image, channel) plane; all threads of the CTA call it.
template <typename T>
__device__ __forceinline__ void rmq_build_plane(const T* __restrict__ src, float* plane, unsigned short* tabs,
                                                int hw, int W, unsigned magic_w) {
#define TAB(s) (tabs + (size_t)(s) * hw)
  for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
    plane[p] = Elem<T>::to_f(src[p]);
    TAB(0)[p] = (unsigned short)p;  // identity level (0,0)
  }
  __syncthreads();
  rmq_build_level<false>(plane, TAB(0), TAB(2), hw, W, magic_w, 1);  // (0,1)
  rmq_build_level<true>(plane, TAB(0), TAB(1), hw, W, magic_w, 1);   // (1,0)
  __syncthreads();
  rmq_build_level<true>(plane, TAB(2), TAB(3), hw, W, magic_w, 1);   // (1,1)
  __syncthreads();
  rmq_build_level<false>(plane, TAB(3), TAB(5), hw, W, magic_w, 2);  // (1,2)
  rmq_build_level<true>(plane, TAB(3), TAB(4), hw, W, magic_w, 2);   // (2,1)
  __syncthreads();
  rmq_build_level<true>(plane, TAB(5), TAB(6), hw, W, magic_w, 2);   // (2,2)
  __syncthreads();
  rmq_build_level<false>(plane, TAB(6), TAB(8), hw, W, magic_w, 4);  // (2,3)
  rmq_build_level<true>(plane, TAB(6), TAB(7), hw, W, magic_w, 4);   // (3,2)
  __syncthreads();
  rmq_build_level<true>(plane, TAB(8), TAB(9), hw, W, magic_w, 4);   // (3,3)
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

// 2 x 2 windows of a stored level: byte offsets a0/a1 (h part) + b0, b0 + db (w part) into the tables.
__device__ __forceinline__ RmqCand rmq_query_fast(const float* plane, const unsigned short* tabs, unsigned a0,
                                                  unsigned a1, unsigned b0, unsigned db) {
  const char* tb = (const char*)tabs;
  const unsigned ia = *(const unsigned short*)(tb + (a0 + b0));
  const unsigned ib = *(const unsigned short*)(tb + (a0 + b0 + db));
  const unsigned ic = *(const unsigned short*)(tb + (a1 + b0));
  const unsigned id = *(const unsigned short*)(tb + (a1 + b0 + db));
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
  const unsigned magic_w = (unsigned)((0x100000000ull + W - 1) / W);
  const unsigned magic_pw = (unsigned)((0x100000000ull + PW - 1) / PW);
  const int PP = PH * PW, wstride = rmq_wstride(PW);
  const int step_r = kRmqThreads / PP, step_b = kRmqThreads - step_r * PP;
  const int rl_init = threadIdx.x / PP, bin_init = threadIdx.x - rl_init * PP;
  const bool ungrouped = *ws.ungrouped != 0;

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
      rmq_build_plane(input + (size_t)nc * hw, plane, tabs, hw, W, magic_w);
    }
    const size_t cpp = (size_t)channels * PP;
    T* out_c = output + (size_t)it.c * PP;
    A* arg_c = argmax + (size_t)it.c * PP;
    const unsigned nkey = (unsigned)it.n & 0xFFFFFFu;
    int rl = rl_init, bin = bin_init;
    while (rl < it.nr) {
      const int r = it.r0 + rl;
      const int ph = fast_div(bin, magic_pw);
      const int pw = bin - ph * PW;
      const uint4 hd = __ldg(ws.hdesc + (size_t)r * PH + ph);
      const unsigned wd = __ldg(ws.wdesc + (size_t)r * wstride + pw);
      const size_t o = (size_t)r * cpp + bin;
      // advance (r, bin) by kRmqThreads outputs without a division
      bin += step_b;
      rl += step_r;
      if (bin >= PP) {
        bin -= PP;
        ++rl;
      }
      if (ungrouped && (hd.w & 0xFFFFFFu) != nkey) continue;
      RmqCand best;
      if (hd.w & wd & kRmqGoodMask) {
        best = rmq_query_fast(plane, tabs, hd.x, hd.y, wd & 0xFFFFFu, (wd >> 20) & 0x1Fu);
      } else {
        best = rmq_query_slow(plane, tabs, hw, W, __ldg(ws.hbins + (size_t)r * PH + ph),
                              __ldg(ws.wbins + (size_t)r * PW + pw));
      }
      out_c[o] = Elem<T>::from_f(__fmul_rn(best.v, __uint_as_float(hd.z)));
      arg_c[o] = Arg<A>::enc((int)best.i);
    }
  }
}

}  // namespace jtsm
