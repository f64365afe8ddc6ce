// RoIPool forward for 16-bit feature maps (bf16 / fp16): range-max tables with PACKED (value, index) entries.
//
// A 16-bit value and a 16-bit cell index fit one 32-bit word, so for these element types a table entry can carry the
// window's maximum ITSELF, exactly:
//     entry = okey16(v) << 16 | (0xFFFF - index)
//     okey16: the order-preserving map of the 16-bit float pattern (sign flip), -0 folded onto +0 (they compare equal in
//             the reference's `>` scan); a cell the scan can never select (NaN, -inf) gets key 0, below every real value.
// Unsigned integer maximum == (value desc, index asc): the first arg-max in row-major order, the reference's tie rule
// (ROILoopPool_cpu.cpp:55-66 / torchvision roi_pool).  A bin = 2 x 2 windows = 4 look-ups + 3 integer max and the
// winner's value and index are in the word: no index -> value look-ups (the index-table kernel of roi_pool_rmq.cuh
// needs 8 dependent look-ups and widens the plane to fp32).  The build is 4 loads + 3 max per cell and level triple.
// Price: 40 B per cell (10 levels x 4 B, level (0,0) IS the packed plane) against 24 B, so 75 x 100 maps take two row
// bands (roi_bins.cuh RmqBands); descriptors, schedule, slow list and band logic are shared with roi_pool_rmqk.cuh.
// A pooled -0.0 is reported as +0.0 (equal under ==; the argmax is the reference's).
#pragma once
#include "roi_pool_rmqk.cuh"

namespace jtsm {

constexpr int kRpHeader = 64;  // mbarrier [0,8)
__host__ __device__ inline size_t rp_smem_core(int hw) { return (size_t)kRpHeader + (((size_t)kRmqTables * hw * 4 + 15) & ~(size_t)15); }

template <typename T> struct Bits16;
template <> struct Bits16<__nv_bfloat16> {
  static __device__ __forceinline__ unsigned short get(__nv_bfloat16 v) { return __bfloat16_as_ushort(v); }
  static __device__ __forceinline__ float to_f(unsigned short b) { return __bfloat162float(__ushort_as_bfloat16(b)); }
};
template <> struct Bits16<__half> {
  static __device__ __forceinline__ unsigned short get(__half v) { return __half_as_ushort(v); }
  static __device__ __forceinline__ float to_f(unsigned short b) { return __half2float(__ushort_as_half(b)); }
};
template <> struct Bits16<float> {  // never instantiated for real work (the launcher packs 16-bit types only)
  static __device__ __forceinline__ unsigned short get(float) { return 0; }
  static __device__ __forceinline__ float to_f(unsigned short) { return 0.f; }
};

template <typename T>
__device__ __forceinline__ unsigned rp_pack(T v, unsigned idx) {
  const float x = Elem<T>::to_f(v);
  unsigned b = Bits16<T>::get(v);
  if (b == 0x8000u) b = 0u;                                               // -0 == +0
  const unsigned key = (b & 0x8000u) ? (~b & 0xFFFFu) : (b | 0x8000u);    // monotone in the value
  return (x > -FLT_MAX ? key << 16 : 0u) | (0xFFFFu - idx);              // NaN / -inf: never selected
}
template <typename T>
__device__ __forceinline__ float rp_value(unsigned e) {  // key != 0
  const unsigned k = e >> 16;
  return Bits16<T>::to_f((unsigned short)((k & 0x8000u) ? (k ^ 0x8000u) : (~k & 0xFFFFu)));
}

__device__ __forceinline__ unsigned lds_u32(unsigned addr) {
  unsigned v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

// from level (k, k) in `src` to (k, k+1), (k+1, k), (k+1, k+1); same clamping argument as rmq_build_phase
__device__ __forceinline__ void rp_build_phase(const unsigned* src, unsigned* d_w, unsigned* d_h, unsigned* d_hw, int hw,
                                               int W, int s) {
  const int last = hw - 1;
#pragma unroll 2
  for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
    const unsigned a = src[p], b = src[min(p + s, last)], c = src[min(p + s * W, last)], d = src[min(p + s * W + s, last)];
    const unsigned ab = max(a, b), cd = max(c, d);
    d_w[p] = ab;
    d_h[p] = max(a, c);
    d_hw[p] = max(ab, cd);
  }
}
__device__ __forceinline__ void rp_build_tables(unsigned* tabs, int hw, int W) {
#define TAB(s) (tabs + (size_t)(s) * hw)
  rp_build_phase(TAB(0), TAB(2), TAB(1), TAB(3), hw, W, 1);
  __syncthreads();
  rp_build_phase(TAB(3), TAB(5), TAB(4), TAB(6), hw, W, 2);
  __syncthreads();
  rp_build_phase(TAB(6), TAB(8), TAB(7), TAB(9), hw, W, 4);
  __syncthreads();
#undef TAB
}

// slow bins (empty / elongated / long ranges): nearest stored level, as many windows as needed.  Empty: 0xFFFFFFFF flag.
struct RpCand {
  unsigned e;
  bool empty;
};
static __device__ __noinline__ RpCand rp_query_slow(const unsigned* tabs, int hw, int W, uint32_t hb, uint32_t wb) {
  const int hs = range_lo(hb), he = range_hi(hb), wst = range_lo(wb), wen = range_hi(wb);
  const int lh = he - hs, lw = wen - wst;
  if (lh <= 0 || lw <= 0) return {0u, true};
  const int qh = min(kRmqMaxLevel, 31 - __clz(lh));
  const int qw = min(kRmqMaxLevel, 31 - __clz(lw));
  const int kh = min(qh, qw + 1), kw = min(qw, qh + 1);
  const unsigned* tab = tabs + (size_t)rmq_slot(kh, kw) * hw;
  const int sh = 1 << kh, sw = 1 << kw;
  const int hl = he - sh, wl = wen - sw;
  unsigned best = 0u;
  for (int hh = hs;; hh += sh) {
    const int h2 = min(hh, hl);
    for (int ww = wst;; ww += sw) {
      const int w2 = min(ww, wl);
      best = max(best, tab[h2 * W + w2]);
      if (ww >= wl) break;
    }
    if (hh >= hl) break;
  }
  return {best, false};
}

template <typename T, typename A, bool kUngrouped, bool kBanded>
__device__ __forceinline__ void rp_main_loop(const unsigned* tabs, const PoolWorkspace& ws, const RmqItem& it, int PP,
                                             int cpp, unsigned W, unsigned magic_pp, int band, unsigned arg_off,
                                             T* __restrict__ output, A* __restrict__ argmax) {
  const unsigned i_end = (unsigned)((it.r0 + it.nr) * PP);
  const unsigned o_c = (unsigned)(it.c * PP), o_r = (unsigned)(cpp - PP);
  const unsigned tabs_s = smem_u32(tabs);
  constexpr int K = kRmqBinsPerThread;
  unsigned i = (unsigned)(it.r0 * PP) + threadIdx.x;
  unsigned xn[K];
  float scn[K];
  auto fetch = [&](unsigned i0) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const unsigned ik = i0 + k * kRmqThreads;
      xn[k] = 0u, scn[k] = 0.f;
      if (ik < i_end) {
        xn[k] = __ldg(ws.bdesc + ik);
        if (kBanded && (int)__ldg(ws.bband + ik) != band) xn[k] = 0u;
        scn[k] = __ldg(ws.rscale + __umulhi(ik, magic_pp));
      }
    }
  };
  fetch(i);
  while (i < i_end) {
    unsigned x[K], o[K];
    float sc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const unsigned ik = i + k * kRmqThreads;
      x[k] = xn[k], sc[k] = scn[k];
      o[k] = ik + __umulhi(ik, magic_pp) * o_r + o_c;
    }
    i += K * kRmqThreads;
    fetch(i);
    if (kBanded) {
      bool any = false;
#pragma unroll
      for (int k = 0; k < K; ++k) any |= (int)x[k] < 0;
      if (!__any_sync(__activemask(), any)) continue;
    }
    unsigned best[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      // descriptor fields are byte offsets of 16-bit tables (roi_bins.cuh): doubled for 32-bit entries
      const unsigned base = tabs_s + 2u * (x[k] & 0x3FFFFu), dh = ((x[k] >> 18) & 15u) * (4u * W), db = 2u * ((x[k] >> 22) & 31u);
      best[k] = max(max(lds_u32(base), lds_u32(base + db)), max(lds_u32(base + dh), lds_u32(base + dh + db)));
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      bool act = (int)x[k] < 0;
      if (kUngrouped && act) act = __ldg(ws.batch_idx + (o[k] - o_c) / (unsigned)cpp) == it.n;
      if (act) {
        const bool none = (best[k] >> 16) == 0u;  // nothing beats the scan's initial -FLT_MAX
        const float v = none ? -FLT_MAX : rp_value<T>(best[k]);
        output[o[k]] = Elem<T>::from_f(__fmul_rn(v, sc[k]));
        const unsigned idx = 0xFFFFu - (best[k] & 0xFFFFu);
        argmax[o[k]] = none ? Arg<A>::enc(-1) : (A)(kBanded ? idx + arg_off : idx);
      }
    }
  }
}

// banded plane over its band's segment of the bin-id list (see rk_main_loop_list): positions [j0, j1)
template <typename T, typename A>
__device__ __forceinline__ void rp_main_loop_list(const unsigned* tabs, const PoolWorkspace& ws, int c, int j0, int j1, int PP,
                                                  int cpp, unsigned W, unsigned magic_pp, unsigned arg_off,
                                                  T* __restrict__ output, A* __restrict__ argmax) {
  const unsigned o_c = (unsigned)(c * PP), o_r = (unsigned)(cpp - PP);
  const unsigned tabs_s = smem_u32(tabs);
  constexpr int K = kRmqBinsPerThread;
  unsigned j = (unsigned)j0 + threadIdx.x;
  const unsigned j_end = (unsigned)j1;
  unsigned xn[K], idn[K];
  float scn[K];
  auto fetch = [&](unsigned jj) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const unsigned jk = jj + k * kRmqThreads;
      xn[k] = 0u, idn[k] = 0u, scn[k] = 0.f;
      if (jk < j_end) {
        idn[k] = (unsigned)__ldg(ws.band_list + jk);
        xn[k] = __ldg(ws.bdesc + idn[k]);
        scn[k] = __ldg(ws.rscale + __umulhi(idn[k], magic_pp));
      }
    }
  };
  fetch(j);
  while (j < j_end) {
    unsigned x[K], o[K], best[K];
    float sc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      x[k] = xn[k], sc[k] = scn[k];
      o[k] = idn[k] + __umulhi(idn[k], magic_pp) * o_r + o_c;
    }
    j += K * kRmqThreads;
    fetch(j);
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const unsigned base = tabs_s + 2u * (x[k] & 0x3FFFFu), dh = ((x[k] >> 18) & 15u) * (4u * W), db = 2u * ((x[k] >> 22) & 31u);
      best[k] = max(max(lds_u32(base), lds_u32(base + db)), max(lds_u32(base + dh), lds_u32(base + dh + db)));
    }
#pragma unroll
    for (int k = 0; k < K; ++k)
      if ((int)x[k] < 0) {
        const bool none = (best[k] >> 16) == 0u;
        const float v = none ? -FLT_MAX : rp_value<T>(best[k]);
        output[o[k]] = Elem<T>::from_f(__fmul_rn(v, sc[k]));
        argmax[o[k]] = none ? Arg<A>::enc(-1) : (A)(0xFFFFu - (best[k] & 0xFFFFu) + arg_off);
      }
  }
}

template <typename T, typename A>
__device__ __forceinline__ void rp_slow_pass(const unsigned* tabs, const PoolWorkspace& ws, const T* __restrict__ input,
                                             const RmqVPlane& vpl, int r_lo, int r_hi, bool ungrouped, int channels, int hw,
                                             int H, int W, int PH, int PW, const RmqBands& bands, T* __restrict__ output,
                                             A* __restrict__ argmax, int chunks = 0) {
  const int PP = PH * PW;  // chunks > 0: band-list mode, [r_lo, r_hi) are chunk numbers (see rk_slow_pass)
  const int n = vpl.nc / channels, c = vpl.nc - n * channels;
  const int ns = __ldg(ws.slow_count);
  for (int j = threadIdx.x; j < ns; j += kRmqThreads) {
    const int e = __ldg(ws.slow_list + j);
    const int id = e & 0x3FFFFFFF;
    const int rr = id / PP, b = id - rr * PP;
    if (chunks > 0) {
      const int jc = j % chunks;
      if (jc < r_lo || jc >= r_hi || __ldg(ws.batch_idx + rr) != n) continue;
    } else {
      if (rr < r_lo || rr >= r_hi) continue;
      if (ungrouped && __ldg(ws.batch_idx + rr) != n) continue;
    }
    const unsigned so = ((unsigned)rr * (unsigned)channels + (unsigned)c) * (unsigned)PP + (unsigned)b;
    const int ph = b / PW, pw = b - ph * PW;
    const uint32_t hb = __ldg(ws.hbins + (size_t)rr * PH + ph), wb = __ldg(ws.wbins + (size_t)rr * PW + pw);
    float v;
    int idx;
    if (e & 0x40000000) {  // taller than a band: scanned from global memory, once, by band 0's CTA
      if (vpl.band != 0) continue;
      const T* src = input + (size_t)vpl.nc * H * W;
      v = -FLT_MAX, idx = -1;
      for (int h = range_lo(hb); h < range_hi(hb); ++h)
        for (int w = range_lo(wb); w < range_hi(wb); ++w) {
          const float x = Elem<T>::to_f(src[h * W + w]);
          if (x > v) v = x, idx = h * W + w;
        }
      if (range_hi(wb) <= range_lo(wb)) v = 0.f;
    } else {
      if (band_of(range_lo(hb), bands) != vpl.band) continue;
      const uint32_t hb_rel = hb - ((uint32_t)vpl.row0 | ((uint32_t)vpl.row0 << 16));
      const RpCand q = rp_query_slow(tabs, hw, W, hb_rel, wb);
      if (q.empty) v = 0.f, idx = -1;
      else if ((q.e >> 16) == 0u) v = -FLT_MAX, idx = -1;
      else v = rp_value<T>(q.e), idx = (int)(0xFFFFu - (q.e & 0xFFFFu)) + vpl.row0 * W;
    }
    output[so] = Elem<T>::from_f(__fmul_rn(v, __ldg(ws.rscale + rr)));
    argmax[so] = Arg<A>::enc(idx);
  }
}

// shared memory: header | 10 tables u32 [hw] | staging (raw 16-bit plane, when it fits)
template <typename T, typename A, bool kBanded>
__global__ void __launch_bounds__(kRmqThreads, 1)
roi_pool_fwd_rmqp_kernel(const T* __restrict__ input, unsigned long long input_bytes, int planes, int channels, int H,
                         int W, int PH, int PW, int chunks, int num_items, int staged, RmqBands bands, PoolWorkspace ws,
                         T* __restrict__ output, A* __restrict__ argmax) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const bool band_list = kBanded && (staged & 4) != 0;  // bit 2 of `staged`: per-band RoI lists
  staged &= 1;
  const int hw = bands.Hb * W;
  unsigned long long* bar = (unsigned long long*)smem_raw;
  unsigned* tabs = (unsigned*)(smem_raw + kRpHeader);
  unsigned char* stage = smem_raw + rp_smem_core(hw);
  const int PP = PH * PW, cpp = channels * PP;
  const bool ungrouped = *ws.ungrouped != 0;
  const unsigned magic_pp = (unsigned)((0x100000000ull + PP - 1) / PP);
  const unsigned a0 = (unsigned)((uintptr_t)input & 15u), in_end = a0 + (unsigned)input_bytes;
  const unsigned char* in_al = (const unsigned char*)input - a0;

  RmqSchedule sch;
  sch.init(num_items, chunks);
  int vp, chunk;
  sch.first(vp, chunk);
  auto window = [&](int pl, unsigned& src_off, unsigned& off, unsigned& bytes) {
    const RmqVPlane v = rmq_vplane(pl, planes, H, W, bands);
    const unsigned start = a0 + (unsigned)(v.elem_off * sizeof(T));
    src_off = start & ~15u;
    off = start - src_off;
    bytes = (off + (unsigned)v.cells * (unsigned)sizeof(T) + 15u) & ~15u;
    return src_off >= a0 && src_off + bytes <= in_end;
  };
  auto issue = [&](int pl) {
    unsigned src_off, off, bytes;
    if (!window(pl, src_off, off, bytes)) return false;
    if (threadIdx.x == 0) {
      fence_proxy_async();
      mbar_expect_tx(bar, bytes);
      bulk_copy_g2s(stage, in_al + src_off, bytes, bar);
    }
    return true;
  };
  if (staged) {
    if (threadIdx.x == 0) {
      mbar_init(bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
  }
  unsigned parity = 0;
  int in_flight = -1;
  if (staged && sch.my_items > 0 && issue(vp)) in_flight = vp;

  int cur_vp = -1;
  RmqVPlane cur{};
  int r_lo = 0, r_hi = 0;
  for (int idx = 0; idx < sch.my_items; ++idx) {
    if (idx > 0) sch.advance(idx, vp, chunk);
    const RmqVPlane v = rmq_vplane(vp, planes, H, W, bands);
    const RmqItem it = rmq_item(v.nc, chunk, chunks, channels, ws);
    if (vp != cur_vp) {
      if (cur_vp >= 0)
        rp_slow_pass<T, A>(tabs, ws, input, cur, r_lo, r_hi, ungrouped, channels, hw, H, W, PH, PW, bands, output, argmax,
                           band_list ? chunks : 0);
      __syncthreads();  // everyone is done with the old tables
      if (in_flight == vp) {
        mbar_wait(bar, parity);
        parity ^= 1u;
        unsigned src_off, off, bytes;
        window(vp, src_off, off, bytes);
        const T* sp = (const T*)(stage + off);
        for (int p = threadIdx.x; p < v.cells; p += kRmqThreads) tabs[p] = rp_pack<T>(sp[p], (unsigned)p);
      } else {
        const T* src = input + v.elem_off;
        for (int p = threadIdx.x; p < v.cells; p += kRmqThreads) tabs[p] = rp_pack<T>(src[p], (unsigned)p);
      }
      if (kBanded)
        for (int p = v.cells + threadIdx.x; p < hw; p += kRmqThreads) tabs[p] = 0xFFFFu - (unsigned)p;  // shorter last band
      __syncthreads();  // level (0,0) complete, staging buffer free
      in_flight = -1;
      if (staged) {
        const int nx = sch.next_plane(idx, vp, chunk);
        if (nx >= 0 && issue(nx)) in_flight = nx;
      }
      cur_vp = vp;
      cur = v;
      r_lo = it.r0, r_hi = it.r0;
      if (band_list) r_lo = chunk, r_hi = chunk;
      rp_build_tables(tabs, hw, W);
    }
    if (it.nr > 0) r_hi = it.r0 + it.nr;
    if (band_list) {
      r_hi = chunk + 1;
      const int slot = it.n * kBandListBands + v.band;
      const int off = __ldg(ws.band_off + slot), cnt = __ldg(ws.band_off + slot + 1) - off;
      const int j0 = off + (int)((long long)cnt * chunk / chunks), j1 = off + (int)((long long)cnt * (chunk + 1) / chunks);
      rp_main_loop_list<T, A>(tabs, ws, it.c, j0, j1, PP, cpp, (unsigned)W, magic_pp, (unsigned)(v.row0 * W), output, argmax);
    } else if (ungrouped)
      rp_main_loop<T, A, true, kBanded>(tabs, ws, it, PP, cpp, (unsigned)W, magic_pp, v.band, (unsigned)(v.row0 * W), output,
                                        argmax);
    else
      rp_main_loop<T, A, false, kBanded>(tabs, ws, it, PP, cpp, (unsigned)W, magic_pp, v.band, (unsigned)(v.row0 * W),
                                         output, argmax);
  }
  if (cur_vp >= 0)
    rp_slow_pass<T, A>(tabs, ws, input, cur, r_lo, r_hi, ungrouped, channels, hw, H, W, PH, PW, bands, output, argmax,
                       band_list ? chunks : 0);
}

}  // namespace jtsm
