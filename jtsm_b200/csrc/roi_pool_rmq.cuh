// RoIPool forward with per-plane 2-D range-max index tables ("rmq" kernel).
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
// One persistent CTA per SM walks a contiguous range of (plane, RoI-chunk) items, rebuilding the
// tables only when the plane changes (~3.5 builds per CTA at config 2).
#pragma once
#include "roi_bins.cuh"

namespace jtsm {

constexpr int kRmqThreads = 1024;
constexpr int kRmqTables = 10;   // slot = 3*min(kh,kw) + (kh-kw+1); slot 1 = identity, (3,3) -> 9
constexpr int kRmqMaxLevel = 3;  // windows of 1,2,4,8 cells per axis

__host__ __device__ inline size_t rmq_smem_bytes(int hw) {
  return (size_t)hw * 4 + (size_t)kRmqTables * hw * 2 + 16;
}

__device__ __forceinline__ int rmq_slot(int kh, int kw) {
  const int m = min(kh, kw);
  return m == kRmqMaxLevel ? 9 : 3 * m + (kh - kw + 1);
}

// exact h = p / W for p < 65536 (magic = ceil(2^32 / W))
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

// Slow path of one bin (elongated or > 16-cell ranges): nearest stored level, as many windows as needed.
__device__ __noinline__ RmqCand rmq_query_generic(const float* plane, const unsigned short* tabs, int hw, int W,
                                                  int hs, int lh, int wst, int lw) {
  const int he = hs + lh, wen = wst + lw;
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

template <typename T, typename A>
__global__ void __launch_bounds__(kRmqThreads, 1)
roi_pool_fwd_rmq_kernel(const T* __restrict__ input, int channels, int H, int W, int PH, int PW,
                        int chunks, int num_items, PoolWorkspace ws, T* __restrict__ output,
                        A* __restrict__ argmax) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int hw = H * W;
  float* plane = (float*)smem_raw;
  unsigned short* tabs = (unsigned short*)(plane + hw);
#define TAB(s) (tabs + (size_t)(s) * hw)
  const unsigned magic_w = (unsigned)((0x100000000ull + W - 1) / W);
  const unsigned magic_pw = (unsigned)((0x100000000ull + PW - 1) / PW);
  const int PP = PH * PW;
  const int step_r = kRmqThreads / PP, step_b = kRmqThreads - step_r * PP;
  const int rl_init = threadIdx.x / PP, bin_init = threadIdx.x - rl_init * PP;
  const bool ungrouped = *ws.ungrouped != 0;

  // contiguous item range of this CTA
  const int per = num_items / gridDim.x, rem = num_items % gridDim.x;
  const int i0 = blockIdx.x * per + min((int)blockIdx.x, rem);
  const int i1 = i0 + per + ((int)blockIdx.x < rem ? 1 : 0);
  int cur_plane = -1;

  for (int item = i0; item < i1; ++item) {
    const int nc = item / chunks;
    const int chunk = item - nc * chunks;
    const int n = nc / channels;
    const int c = nc - n * channels;
    if (nc != cur_plane) {
      __syncthreads();  // previous item's queries are done with the old tables
      cur_plane = nc;
      const T* src = input + (size_t)nc * hw;
      for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
        plane[p] = Elem<T>::to_f(src[p]);
        TAB(1)[p] = (unsigned short)p;  // identity level (0,0)
      }
      __syncthreads();
      rmq_build_level<false>(plane, TAB(1), TAB(0), hw, W, magic_w, 1);  // (0,1)
      rmq_build_level<true>(plane, TAB(1), TAB(2), hw, W, magic_w, 1);   // (1,0)
      __syncthreads();
      rmq_build_level<true>(plane, TAB(0), TAB(4), hw, W, magic_w, 1);   // (1,1)
      __syncthreads();
      rmq_build_level<false>(plane, TAB(4), TAB(3), hw, W, magic_w, 2);  // (1,2)
      rmq_build_level<true>(plane, TAB(4), TAB(5), hw, W, magic_w, 2);   // (2,1)
      __syncthreads();
      rmq_build_level<true>(plane, TAB(3), TAB(7), hw, W, magic_w, 2);   // (2,2)
      __syncthreads();
      rmq_build_level<false>(plane, TAB(7), TAB(6), hw, W, magic_w, 4);  // (2,3)
      rmq_build_level<true>(plane, TAB(7), TAB(8), hw, W, magic_w, 4);   // (3,2)
      __syncthreads();
      rmq_build_level<true>(plane, TAB(6), TAB(9), hw, W, magic_w, 4);   // (3,3)
      __syncthreads();
    }
    const int first = ws.first[n];
    const int last = -ws.neg_last[n];
    if (first > last) continue;
    const int len = last - first + 1;
    const int r0 = first + (int)(((long long)len * chunk) / chunks);
    const int nr = first + (int)(((long long)len * (chunk + 1)) / chunks) - r0;
    const size_t cpp = (size_t)channels * PP;
    T* out_c = output + (size_t)c * PP;
    A* arg_c = argmax + (size_t)c * PP;
    int rl = rl_init, bin = bin_init;
    while (rl < nr) {
      const int r = r0 + rl;
      const int ph = fast_div(bin, magic_pw);
      const int pw = bin - ph * PW;
      const uint4 hd = __ldg(ws.hdesc + (size_t)r * PH + ph);
      const uint2 wd = __ldg(ws.wdesc + (size_t)r * PW + pw);
      const size_t o = (size_t)r * cpp + bin;
      // advance (r, bin) by kRmqThreads outputs without a division
      bin += step_b;
      rl += step_r;
      if (bin >= PP) {
        bin -= PP;
        ++rl;
      }
      if (ungrouped && (int)hd.w != n) continue;
      const unsigned fl = hd.y | wd.y;
      float m = 0.f;
      int mi = -1;
      if (!(fl & 8u)) {  // non-empty bin
        const int qh = hd.y & 3, qw = wd.y & 3;
        RmqCand best;
        if (!(fl & 4u) && abs(qh - qw) <= 1) {  // 2 x 2 windows of a stored level
          const unsigned short* tab = TAB(rmq_slot(qh, qw));
          const unsigned h0 = hd.x & 0xFFFFu, h1 = hd.x >> 16, w0 = wd.x & 0xFFFFu, w1 = wd.x >> 16;
          const unsigned ia = tab[h0 + w0], ib = tab[h0 + w1], ic = tab[h1 + w0], id = tab[h1 + w1];
          best = rmq_best(rmq_best(rmq_load(plane, ia), rmq_load(plane, ib)),
                          rmq_best(rmq_load(plane, ic), rmq_load(plane, id)));
        } else {
          const int hs = fast_div((int)(hd.x & 0xFFFFu), magic_w);
          best = rmq_query_generic(plane, tabs, hw, W, hs, (int)(hd.y >> 8), (int)(wd.x & 0xFFFFu), (int)(wd.y >> 8));
        }
        m = best.v;
        mi = (int)best.i;
      }
      m = __fmul_rn(m, __uint_as_float(hd.z));
      out_c[o] = Elem<T>::from_f(m);
      arg_c[o] = Arg<A>::enc(mi);
    }
  }
#undef TAB
}

}  // namespace jtsm
