// RoIPool forward, range-max tables with KEYED entries ("rmqk" kernel) -- the round-2 form of roi_pool_rmq.cuh.
//
// roi_pool_rmq.cuh answers a bin with 4 index look-ups + 4 value look-ups (each index must be turned into a value
// before two candidates can be compared); the kernel is bound by the shared-memory data pipe (81 % of peak,
// profiles/r1u), so the only way down is fewer look-ups.  Here a table entry carries enough of the VALUE to decide
// almost every comparison without touching the plane:
//
//   entry (16 bit) = key10 << 6 | dr << 3 | dc
//     key10  a per-plane monotone 10-bit quantisation of the window's maximum: 0 iff the value equals the plane's
//            minimum, else 1 + min(1022, int((v - lo) * 1022 / (hi - lo)));  v1 > v2  =>  key(v1) >= key(v2)
//     dr,dc  position of the window's FIRST maximum relative to the window's origin (windows are <= 8 x 8)
//
// Tables are still built with exact (value desc, index asc) comparisons, so every entry names the true first arg-max
// of its window and the key of the true maximum.  A bin = 4 windows = 4 entries -> 4 packed candidates
// (key << 22 | ~absolute index); their integer maximum names the winner unless another candidate has the SAME key but
// a different cell and the key is not 0 (key 0: all such values equal the plane minimum exactly, so the lowest index
// wins, which is what the integer maximum picks).  That ambiguous case (~1 % of the bins on the synthetic maps) takes
// the exact path -- 4 value look-ups, as roi_pool_rmq.cuh always does.  So: 4 entry look-ups + 1 value look-up per
// bin instead of 8, same 16-bit tables, same shared-memory footprint, results bit-identical to the scan
// (ROILoopPool_cpu.cpp:55-66) by construction: keys only ever decide comparisons they are able to decide.
//
// Non-finite cells: a cell that can never win the reference's strict `>` scan from -FLT_MAX (NaN, -inf, -FLT_MAX)
// is staged as -FLT_MAX, and a bin whose maximum is -FLT_MAX reports (-FLT_MAX, -1) exactly like the scan; +inf
// makes the key scale 0 (every key 1 -> exact path), slow but right.
//
// Plane fetch: one `cp.async.bulk` (TMA, 1-D) + mbarrier per plane into a raw staging buffer -- issued by one thread
// as soon as the previous plane has been unpacked, it lands behind the table build and the queries of the current
// plane and costs no LSU wavefronts (the cp.async form of roi_pool_rmq.cuh: 7500 4-byte LDGSTS per plane, fp32 only).
// Any element type: the unpack pass (convert, sanitise, min / max) reads the staged bytes.  16-byte alignment is
// reached by fetching the aligned window around the plane (<= 15 bytes of the neighbouring planes on either side);
// a window that would pass the end of the tensor falls back to plain loads.
#pragma once
#include "roi_pool_rmq.cuh"

namespace jtsm {

// ---- mbarrier / bulk-copy primitives (PTX; SASS: SYNCS.*, UBLKCP) -------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
      "selp.u32 %0, 1, 0, P1;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// try_wait suspends the thread for a bounded time per call; a copy that never lands (a bug, not a state the kernel
// can be in) traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  for (unsigned it = 0; !mbar_try_wait(bar, parity); ++it)
    if (it > (1u << 24)) __trap();
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- keyed entries -------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned rk_abs(unsigned e, unsigned pos, unsigned W) {
  return pos + ((e >> 3) & 7u) * W + (e & 7u);
}
__device__ __forceinline__ unsigned rk_key(float v, float lo, float scale) {
  const int q = (int)__fmul_rn(__fsub_rn(v, lo), scale);  // monotone in v; the conversion saturates, NaN -> 0
  return v == lo ? 0u : (unsigned)(1 + min(q, 1022));
}
__device__ __forceinline__ bool rk_better(float vb, unsigned ib, float va, unsigned ia) {
  return (vb > va) || (vb == va && ib < ia);
}

// One build phase (see rmq_build_phase): level (k, k) in `src` -> (k, k+1) wide, (k+1, k) tall, (k+1, k+1).
// The winner's entry is inherited: its key is the key of the new window's maximum, its (dr, dc) are shifted by the
// child window's offset (dc + s <= 7 and dr + s <= 7, so the fields never carry into each other).  Entries whose
// window leaves the plane are never queried; their decoded index is clamped so the build never reads past the plane.
template <bool kIdentity>
__device__ __forceinline__ void rk_build_phase(const float* plane, const unsigned short* src, unsigned short* d0,
                                               unsigned short* d_w, unsigned short* d_h, unsigned short* d_hw, int hw,
                                               int W, int s, float lo, float scale) {
  const unsigned last = (unsigned)(hw - 1);
#pragma unroll 2
  for (int p = threadIdx.x; p < hw; p += kRmqThreads) {
    const unsigned pa = (unsigned)p, pb = min(pa + s, last), pc = min(pa + s * W, last), pd = min(pa + s * W + s, last);
    unsigned ea, eb, ec, ed, ia, ib, ic, id;
    if (kIdentity) {
      ia = pa, ib = pb, ic = pc, id = pd;
    } else {
      ea = src[pa], eb = src[pb], ec = src[pc], ed = src[pd];
      ia = min(rk_abs(ea, pa, W), last), ib = min(rk_abs(eb, pb, W), last);
      ic = min(rk_abs(ec, pc, W), last), id = min(rk_abs(ed, pd, W), last);
    }
    const float va = plane[ia], vb = plane[ib], vc = plane[ic], vd = plane[id];
    if (kIdentity) {
      ea = rk_key(va, lo, scale) << 6, eb = rk_key(vb, lo, scale) << 6;
      ec = rk_key(vc, lo, scale) << 6, ed = rk_key(vd, lo, scale) << 6;
      d0[p] = (unsigned short)ea;  // level (0, 0): the cell itself
    }
    eb += (unsigned)s, ec += (unsigned)s << 3, ed += (unsigned)s + ((unsigned)s << 3);
    const bool b_ab = rk_better(vb, ib, va, ia), b_ac = rk_better(vc, ic, va, ia), b_cd = rk_better(vd, id, vc, ic);
    const float vab = b_ab ? vb : va, vcd = b_cd ? vd : vc;
    const unsigned iab = b_ab ? ib : ia, icd = b_cd ? id : ic;
    const unsigned eab = b_ab ? eb : ea, ecd = b_cd ? ed : ec;
    d_w[p] = (unsigned short)eab;
    d_h[p] = (unsigned short)(b_ac ? ec : ea);
    d_hw[p] = (unsigned short)(rk_better(vcd, icd, vab, iab) ? ecd : eab);
  }
}

__device__ __forceinline__ void rk_build_tables(const float* plane, unsigned short* tabs, int hw, int W, float lo,
                                                float scale) {
#define TAB(s) (tabs + (size_t)(s) * hw)
  rk_build_phase<true>(plane, TAB(0), TAB(0), TAB(2), TAB(1), TAB(3), hw, W, 1, lo, scale);
  __syncthreads();
  rk_build_phase<false>(plane, TAB(3), nullptr, TAB(5), TAB(4), TAB(6), hw, W, 2, lo, scale);
  __syncthreads();
  rk_build_phase<false>(plane, TAB(6), nullptr, TAB(8), TAB(7), TAB(9), hw, W, 4, lo, scale);
  __syncthreads();
#undef TAB
}

// A fast bin: descriptor x (roi_bins.cuh) -> (value, absolute index).
__device__ __forceinline__ RmqCand rk_query_fast(unsigned plane_s, unsigned tabs_s, unsigned x, unsigned W, unsigned hw) {
  const unsigned base = x & 0x3FFFFu, dhc = (x >> 18) & 15u, db = (x >> 22) & 31u, slot = (x >> 27) & 15u;
  const unsigned addr = tabs_s + base, dh = dhc * (2u * W);
  const unsigned ea = lds_u16(addr), eb = lds_u16(addr + db), ec = lds_u16(addr + dh), ed = lds_u16(addr + dh + db);
  const unsigned pa = (base >> 1) - slot * hw, pb = pa + (db >> 1), pc = pa + dhc * W, pd = pc + (db >> 1);
  const unsigned ia = rk_abs(ea, pa, W), ib = rk_abs(eb, pb, W), ic = rk_abs(ec, pc, W), id = rk_abs(ed, pd, W);
  const unsigned ca = ((ea & 0xFFC0u) << 16) | (ia ^ 0xFFFFu), cb = ((eb & 0xFFC0u) << 16) | (ib ^ 0xFFFFu);
  const unsigned cc = ((ec & 0xFFC0u) << 16) | (ic ^ 0xFFFFu), cd = ((ed & 0xFFC0u) << 16) | (id ^ 0xFFFFu);
  const unsigned m = max(max(ca, cb), max(cc, cd));
  RmqCand best;
  best.i = (m & 0xFFFFu) ^ 0xFFFFu;
  best.v = lds_f32(plane_s + 4u * best.i);
  // same key, other cell (x ^ m in [1, 0xFFFF]) and the key is not the plane-minimum key
  const bool amb = (((ca ^ m) - 1u < 0xFFFFu) | ((cb ^ m) - 1u < 0xFFFFu) | ((cc ^ m) - 1u < 0xFFFFu) |
                    ((cd ^ m) - 1u < 0xFFFFu)) && (m >> 22) != 0u;
  if (amb) {  // exact (value desc, index asc) comparison of the four window maxima
    const RmqCand a{lds_f32(plane_s + 4u * ia), ia}, b{lds_f32(plane_s + 4u * ib), ib};
    const RmqCand c{lds_f32(plane_s + 4u * ic), ic}, d{lds_f32(plane_s + 4u * id), id};
    best = rmq_best(rmq_best(a, b), rmq_best(c, d));
  }
  return best;
}

// Slow path of one bin (empty, elongated or > 16-cell ranges): nearest stored level, as many windows as needed,
// exact comparisons.  Empty bin -> value 0, index 0xFFFFFFFF.
static __device__ __noinline__ RmqCand rk_query_slow(const float* plane, const unsigned short* tabs, int hw, int W,
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
  RmqCand best{-FLT_MAX, 0xFFFFFFFFu};
  for (int hh = hs;; hh += sh) {
    const int h2 = min(hh, hl);
    for (int ww = wst;; ww += sw) {
      const int w2 = min(ww, wl);
      const unsigned pos = (unsigned)(h2 * W + w2);
      const unsigned i = rk_abs(tab[pos], pos, (unsigned)W);
      const RmqCand c{plane[i], i};
      best = (best.i == 0xFFFFFFFFu) ? c : rmq_best(best, c);
      if (ww >= wl) break;
    }
    if (hh >= hl) break;
  }
  return best;
}

// ---- item schedule (same as roi_pool_fwd_rmq_kernel: planes round-robin over the CTAs, the partial last round cut into
// chunks) factored so that both the walk and the "which plane comes next" look-ahead use one definition -----------
struct RmqSchedule {
  int G, chunks, full, tail_base, t0, my_items;
  __device__ __forceinline__ void init(int num_items, int chunks_) {
    G = gridDim.x, chunks = chunks_;
    const int planes = num_items / chunks;
    full = (planes / G) * chunks;
    const int tail_items = (planes % G) * chunks;
    tail_base = (planes / G) * G * chunks;
    const int t_per = tail_items / G, t_rem = tail_items % G;
    t0 = blockIdx.x * t_per + min((int)blockIdx.x, t_rem);
    my_items = full + t_per + ((int)blockIdx.x < t_rem ? 1 : 0);
  }
  __device__ __forceinline__ void first(int& pl, int& ch) const {
    pl = full > 0 ? (int)blockIdx.x : (tail_base + t0) / chunks;
    ch = full > 0 ? 0 : (tail_base + t0) % chunks;
  }
  // (pl, ch) of item idx_next - 1 -> item idx_next
  __device__ __forceinline__ void advance(int idx_next, int& pl, int& ch) const {
    if (idx_next == full) {
      pl = (tail_base + t0) / chunks;
      ch = (tail_base + t0) % chunks;
    } else if (++ch == chunks) {
      ch = 0;
      pl += idx_next < full ? G : 1;
    }
  }
  // the plane this CTA serves after `pl` (item idx is the current one), or -1
  __device__ __forceinline__ int next_plane(int idx, int pl, int ch) const {
    int j = idx + 1, p2 = pl, c2 = ch;
    while (j < my_items) {
      advance(j, p2, c2);
      if (p2 != pl) return p2;
      ++j;
    }
    return -1;
  }
};

// One (band, plane) of the schedule: "virtual plane" vp = band * planes + plane (band-major, so that neighbouring CTAs
// serve neighbouring channels of the same band and their output runs share sectors while they are written).
struct RmqVPlane {
  int nc, band, row0, cells;  // (image, channel) plane, band, its first row, its cell count
  size_t elem_off;            // offset of the band's first cell in the input tensor (elements)
};
__device__ __forceinline__ RmqVPlane rmq_vplane(int vp, int planes, int H, int W, const RmqBands& bd) {
  RmqVPlane v;
  v.band = vp / planes;
  v.nc = vp - v.band * planes;
  v.row0 = v.band * bd.S;
  v.cells = band_rows(v.band, H, bd) * W;
  v.elem_off = (size_t)v.nc * H * W + (size_t)v.row0 * W;
  return v;
}

// Main loop of one (plane, RoI-chunk) item: one fast bin per thread and chain, kRmqBinsPerThread independent chains.
// Everything is indexed with 32-bit offsets from the kernel's own pointer arguments (the launcher checks the sizes):
// bin id i -> RoI r = i / PP (magic multiply) -> output element i + r * (C - 1) * PP + c * PP.  Descriptor and row
// scale are loaded one round ahead.  x = 0 marks a slow bin (left to the list) or an inactive thread: the branch-free
// look-ups then run on offset 0, which is harmless.  kKeyed: table format (keyed entries / absolute indices).
template <typename T, typename A, bool kUngrouped, bool kMoi, bool kKeyed, bool kBanded>
__device__ __forceinline__ void rk_main_loop(const float* plane, const unsigned short* tabs, const PoolWorkspace& ws,
                                             const RmqItem& it, int PP, int cpp, unsigned W, unsigned hw,
                                             unsigned magic_pp, const float* __restrict__ mask_scale, int zero_floor,
                                             int nb, int band, unsigned arg_off, T* __restrict__ output,
                                             A* __restrict__ argmax) {
  // kBanded (plane cut into row bands): only the bins of row band `band` are served (ws.bband), their table indices
  // are band relative (arg_off = first cell of the band); warps without a bin of this band skip the look-ups.  The
  // single-band instantiation compiles to the code without any of it.
  const unsigned i_end = (unsigned)((it.r0 + it.nr) * PP);
  const unsigned o_c = (unsigned)(it.c * PP), o_r = (unsigned)(cpp - PP);
  const unsigned plane_s = smem_u32(plane), tabs_s = smem_u32(tabs);
  constexpr int K = kRmqBinsPerThread;
  unsigned i = (unsigned)(it.r0 * PP) + threadIdx.x;
  unsigned xn[K];
  float scn[K], msn[K];
  auto fetch = [&](unsigned i0) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const unsigned ik = i0 + k * kRmqThreads;
      xn[k] = 0u, scn[k] = 0.f, msn[k] = 0.f;
      if (ik < i_end) {
        xn[k] = __ldg(ws.bdesc + ik);
        if (kBanded && (int)__ldg(ws.bband + ik) != band) xn[k] = 0u;  // another band's bin
        scn[k] = __ldg(ws.rscale + __umulhi(ik, magic_pp));
        if (kMoi) msn[k] = __ldg(mask_scale + __umulhi(ik, magic_pp));
      }
    }
  };
  fetch(i);
  while (i < i_end) {
    unsigned x[K], o[K];
    float sc[K], ms[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const unsigned ik = i + k * kRmqThreads;
      x[k] = xn[k], sc[k] = scn[k], ms[k] = msn[k];
      o[k] = ik + __umulhi(ik, magic_pp) * o_r + o_c;
    }
    i += K * kRmqThreads;
    fetch(i);
    if (kBanded) {  // uniform per warp: nothing of this band in the round
      bool any = false;
#pragma unroll
      for (int k = 0; k < K; ++k) any |= (int)x[k] < 0;
      if (!__any_sync(__activemask(), any)) continue;
    }
    RmqCand best[K];
#pragma unroll
    for (int k = 0; k < K; ++k)
      best[k] = kKeyed ? rk_query_fast(plane_s, tabs_s, x[k], W, hw)
                       : rmq_query_fast(plane_s, tabs_s + (x[k] & 0x3FFFFu), ((x[k] >> 18) & 15u) * (2u * W),
                                        (x[k] >> 22) & 31u);
#pragma unroll
    for (int k = 0; k < K; ++k) {
      bool act = (int)x[k] < 0;  // x >= 0: slow bin (left to the list) or past the end
      if (kUngrouped && act) act = __ldg(ws.batch_idx + (o[k] - o_c) / (unsigned)cpp) == it.n;
      if (act) {
        const float v = kMoi ? __fmul_rn(best[k].v, ms[k]) : best[k].v;  // (f * mask_scale) * (objectness + 1)
        // nothing beats the scan's initial maximum: -FLT_MAX (RoIPool, MOIPool) or 0 (ROILoopPool, zero_floor)
        const bool floor0 = kMoi && zero_floor;
        const bool none = floor0 ? !(best[k].v > 0.f) : !(best[k].v > -FLT_MAX);
        output[o[k]] = Elem<T>::from_f(floor0 && none ? 0.f : __fmul_rn(v, sc[k]));
        argmax[o[k]] = none ? Arg<A>::enc(-1) : (A)(kBanded ? best[k].i + arg_off : best[k].i);
      }
    }
  }
}

// Main loop, row-mapped form (index tables, RoIs grouped by image, one band, no mask): the kernels are ISSUE bound
// (profiles/r2o_roi_pool_pack16.md: 75 instructions per bin, a quarter of them look-ups), so this form removes
// bookkeeping instead of look-ups.  A thread keeps ONE bin position and walks the RoIs: thread t -> RoI slot t / PP, bin
// t % PP (G = 1024 / PP RoIs per pass; PP = 49: 980 of 1024 threads work), so the bin id, the output element and the
// row-scale index advance by constants — no division by PP per bin, no per-bin output offset arithmetic — and the two
// descriptor sets ping-pong between two register sets (no carried moves).  Same look-ups, same results.
template <typename T, typename A>
__device__ __forceinline__ void rk_main_loop_rows(const float* plane, const unsigned short* tabs, const PoolWorkspace& ws,
                                                  const RmqItem& it, int PP, int cpp, unsigned W, unsigned magic_pp,
                                                  T* __restrict__ output, A* __restrict__ argmax) {
  const unsigned G = (unsigned)kRmqThreads / (unsigned)PP;
  const unsigned rl = __umulhi(threadIdx.x, magic_pp), bin = threadIdx.x - rl * (unsigned)PP;
  if (rl >= G) return;  // the 1024 % PP threads without a slot
  unsigned plane_s, tabs_s;  // volatile: keeps the shared-window bases in registers instead of re-deriving them per round
  asm volatile("cvt.u32.u64 %0, %1;" : "=r"(plane_s) : "l"(__cvta_generic_to_shared(plane)));
  asm volatile("cvt.u32.u64 %0, %1;" : "=r"(tabs_s) : "l"(__cvta_generic_to_shared(tabs)));
  const unsigned r_end = (unsigned)(it.r0 + it.nr), W2 = 2u * W;
  unsigned r = (unsigned)it.r0 + rl;
  const uint32_t* pd = ws.bdesc + ((size_t)r * PP + bin);
  const float* ps = ws.rscale + r;
  unsigned o = r * (unsigned)cpp + (unsigned)(it.c * PP) + bin;
  const unsigned dd = G * (unsigned)PP, dO = G * (unsigned)cpp;

  struct Set {
    unsigned x0, x1;
    float s0, s1;
  };
  auto fetch = [&](Set& f, unsigned rr, const uint32_t* d, const float* sp) {
    f.x0 = 0u, f.x1 = 0u, f.s0 = 0.f, f.s1 = 0.f;
    if (rr < r_end) f.x0 = __ldg(d), f.s0 = __ldg(sp);
    if (rr + G < r_end) f.x1 = __ldg(d + dd), f.s1 = __ldg(sp + G);
  };
  auto process = [&](const Set& f, unsigned oo) {
    const RmqCand b0 = rmq_query_fast(plane_s, tabs_s + (f.x0 & 0x3FFFFu), ((f.x0 >> 18) & 15u) * W2, (f.x0 >> 22) & 31u);
    const RmqCand b1 = rmq_query_fast(plane_s, tabs_s + (f.x1 & 0x3FFFFu), ((f.x1 >> 18) & 15u) * W2, (f.x1 >> 22) & 31u);
    const T v0 = Elem<T>::from_f(__fmul_rn(b0.v, f.s0)), v1 = Elem<T>::from_f(__fmul_rn(b1.v, f.s1));
    const A a0 = b0.v > -FLT_MAX ? (A)b0.i : Arg<A>::enc(-1), a1 = b1.v > -FLT_MAX ? (A)b1.i : Arg<A>::enc(-1);
    if ((int)f.x0 < 0) {  // x >= 0: slow bin (left to the list) or past the end
      output[oo] = v0;
      argmax[oo] = a0;
    }
    if ((int)f.x1 < 0) {
      output[oo + dO] = v1;
      argmax[oo + dO] = a1;
    }
  };
  if (r >= r_end) return;
  Set fa, fb;
  fetch(fa, r, pd, ps);
  for (;;) {
    fetch(fb, r + 2u * G, pd + 2u * dd, ps + 2u * G);
    process(fa, o);
    r += 2u * G, pd += 2u * dd, ps += 2u * G, o += 2u * dO;
    if (r >= r_end) break;
    fetch(fa, r + 2u * G, pd + 2u * dd, ps + 2u * G);
    process(fb, o);
    r += 2u * G, pd += 2u * dd, ps += 2u * G, o += 2u * dO;
    if (r >= r_end) break;
  }
}

// Main loop of a BANDED plane over its band's segment of the bin-id list (PoolWorkspace::band_list, built by the prep
// kernels): the generic loop above with one indirection -- position -> bin id -> descriptor -- so every lane of every
// round holds a bin of THIS band (the filter form sweeps all bins of the call per (band, plane) pass and runs mostly
// empty warps; a RoI-per-warp list form kept a third of the lanes busy on 150 x 200 maps).  Positions [j0, j1).
template <typename T, typename A, bool kKeyed>
__device__ __forceinline__ void rk_main_loop_list(const float* plane, const unsigned short* tabs, const PoolWorkspace& ws,
                                                  int c, int j0, int j1, int PP, int cpp, unsigned W, unsigned hw,
                                                  unsigned magic_pp, unsigned arg_off, T* __restrict__ output,
                                                  A* __restrict__ argmax) {
  const unsigned o_c = (unsigned)(c * PP), o_r = (unsigned)(cpp - PP);
  const unsigned plane_s = smem_u32(plane), tabs_s = smem_u32(tabs);
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
    unsigned x[K], o[K];
    float sc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      x[k] = xn[k], sc[k] = scn[k];
      o[k] = idn[k] + __umulhi(idn[k], magic_pp) * o_r + o_c;
    }
    j += K * kRmqThreads;
    fetch(j);
    RmqCand best[K];
#pragma unroll
    for (int k = 0; k < K; ++k)
      best[k] = kKeyed ? rk_query_fast(plane_s, tabs_s, x[k], W, hw)
                       : rmq_query_fast(plane_s, tabs_s + (x[k] & 0x3FFFFu), ((x[k] >> 18) & 15u) * (2u * W),
                                        (x[k] >> 22) & 31u);
#pragma unroll
    for (int k = 0; k < K; ++k)
      if ((int)x[k] < 0) {  // x >= 0: slow bin (left to the list) or past the end
        const bool none = !(best[k].v > -FLT_MAX);
        output[o[k]] = Elem<T>::from_f(__fmul_rn(best[k].v, sc[k]));
        argmax[o[k]] = none ? Arg<A>::enc(-1) : (A)(best[k].i + arg_off);
      }
  }
}

// The few bins without a fast descriptor (empty, elongated, > 16 cells: ~2%) of RoIs [r_lo, r_hi) for the plane
// whose tables are resident.  The prep kernel lists them, so no warp of the main loop diverges into the long
// path; the list is short and every CTA filters it for the RoI range it has just served.
template <typename T, typename A, bool kMoi, bool kKeyed>
__device__ __forceinline__ void rk_slow_pass(const float* plane, const unsigned short* tabs, const PoolWorkspace& ws,
                                             const T* __restrict__ input, const RmqVPlane& vpl, int r_lo, int r_hi,
                                             bool ungrouped, int channels, int hw, int H, int W, int PH, int PW,
                                             const RmqBands& bands, T* __restrict__ output, A* __restrict__ argmax,
                                             int chunks = 0) {
  // chunks > 0 (band-list mode): the CTA has served chunks [r_lo, r_hi) of this (band, plane) -- the slow list is dealt
  // by entry number modulo `chunks`, and the image is checked per entry
  const int PP = PH * PW;
  if (kMoi) return;  // the visit-list kernel answers every bin without a fast descriptor
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
    RmqCand best;
    if (e & 0x40000000) {
      // taller than a band (a RoI reaching far outside the map): scanned from global memory, once, by band 0's CTA
      if (vpl.band != 0) continue;
      const T* src = input + (size_t)vpl.nc * H * W;
      best = {-FLT_MAX, 0xFFFFFFFFu};
      for (int h = range_lo(hb); h < range_hi(hb); ++h)
        for (int w = range_lo(wb); w < range_hi(wb); ++w) {
          const float v = Elem<T>::to_f(src[h * W + w]);
          if (v > best.v) best = {v, (unsigned)(h * W + w)};
        }
      if (range_hi(wb) <= range_lo(wb)) best.v = 0.f;  // empty in w (h is non-empty here)
    } else {
      if (band_of(range_lo(hb), bands) != vpl.band) continue;
      const uint32_t hb_rel = hb - ((uint32_t)vpl.row0 | ((uint32_t)vpl.row0 << 16));
      best = kKeyed ? rk_query_slow(plane, tabs, hw, W, hb_rel, wb) : rmq_query_slow(plane, tabs, hw, W, hb_rel, wb);
      if (best.i != 0xFFFFFFFFu) {
        if (!(best.v > -FLT_MAX)) best.i = 0xFFFFFFFFu;  // non-empty, nothing beats -FLT_MAX
        else best.i += (unsigned)(vpl.row0 * W);
      }
    }
    output[so] = Elem<T>::from_f(__fmul_rn(best.v, __ldg(ws.rscale + rr)));
    argmax[so] = Arg<A>::enc((int)best.i);
  }
}

// shared-memory layout of the keyed kernel (bytes): header | plane fp32 [hw] | 10 tables u16 [hw] | staging (raw plane)
constexpr int kRkHeader = 384;  // mbarrier [0,8), plane stats [16,24), reduction scratch [32,288)
__host__ __device__ inline size_t rk_smem_core(int hw) {
  return (size_t)kRkHeader + (((size_t)hw * 4 + 15) & ~(size_t)15) + (((size_t)kRmqTables * hw * 2 + 15) & ~(size_t)15);
}
__host__ __device__ inline size_t rk_stage_bytes(int hw, int esz) { return (((size_t)hw * esz + 15) & ~(size_t)15) + 32; }

template <typename T, typename A, bool kMoi, bool kKeyed, bool kBanded>
__global__ void __launch_bounds__(kRmqThreads, 1)
roi_pool_fwd_rmq_kernel(const T* __restrict__ input, unsigned long long input_bytes, int planes, int channels, int H,
                         int W, int PH, int PW, int chunks, int num_items, int staged, RmqBands bands,
                         PoolWorkspace ws, MoiAux moi, T* __restrict__ output, A* __restrict__ argmax) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const bool rows_loop = (staged & 2) != 0 && PH * PW <= kRmqThreads;  // bit 1 of `staged`: row-mapped main loop
  const bool band_list = kBanded && (staged & 4) != 0;                 // bit 2: per-band RoI lists (banded planes)
  staged &= 1;
  if (kMoi && !moi_rmq_gate(moi.slow_total, moi.total_bins, moi.gate_pct)) return;  // uniform: the list kernel runs
  const int hw = bands.Hb * W;  // cells of a full band = table stride (the whole plane when nb == 1)
  unsigned long long* bar = (unsigned long long*)smem_raw;
  float* stats = (float*)(smem_raw + 16);      // [0] lo, [1] scale
  float* red = (float*)(smem_raw + 32);        // [32 warps][2] partial min / max
  float* plane = (float*)(smem_raw + kRkHeader);
  unsigned short* tabs = (unsigned short*)(smem_raw + kRkHeader + (((size_t)hw * 4 + 15) & ~(size_t)15));
  unsigned char* stage = smem_raw + rk_smem_core(hw);  // valid only when staged
  const int PP = PH * PW, cpp = channels * PP;
  const bool ungrouped = *ws.ungrouped != 0;
  const unsigned magic_pp = (unsigned)((0x100000000ull + PP - 1) / PP);
  // byte offsets are taken from the 16-byte aligned address below `input` (32-bit: the launcher bounds the tensor size)
  const unsigned a0 = (unsigned)((uintptr_t)input & 15u), in_end = a0 + (unsigned)input_bytes;
  const unsigned char* in_al = (const unsigned char*)input - a0;

  RmqSchedule sch;
  sch.init(num_items, chunks);
  int vp, chunk;
  sch.first(vp, chunk);

  // The 16-byte aligned window around the cells of virtual plane `pl`; false when it would leave the tensor (first /
  // last plane of a tensor whose ends are not 16-byte aligned): read with plain loads then.  Uniform per CTA.
  auto window = [&](int pl, unsigned& src_off, unsigned& off, unsigned& bytes) {
    const RmqVPlane v = rmq_vplane(pl, planes, H, W, bands);
    const unsigned start = a0 + (unsigned)(v.elem_off * sizeof(T));
    src_off = start & ~15u;
    off = start - src_off;
    bytes = (off + (unsigned)v.cells * (unsigned)sizeof(T) + 15u) & ~15u;
    return src_off >= a0 && src_off + bytes <= in_end;  // only bytes of the tensor itself are ever fetched
  };
  auto issue = [&](int pl) {
    unsigned src_off, off, bytes;
    if (!window(pl, src_off, off, bytes)) return false;
    if (threadIdx.x == 0) {
      fence_proxy_async();  // the staging buffer was last read through the generic proxy
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
  int in_flight = -1;  // virtual plane whose bytes are on their way into (or already in) the staging buffer
  if (staged && sch.my_items > 0 && issue(vp)) in_flight = vp;

  int cur_vp = -1;
  RmqVPlane cur{};
  int r_lo = 0, r_hi = 0;
  if (!kKeyed)  // index tables: the identity level (0,0), slot 0, never changes
    for (int p = threadIdx.x; p < hw; p += kRmqThreads) tabs[p] = (unsigned short)p;

  for (int idx = 0; idx < sch.my_items; ++idx) {
    if (idx > 0) sch.advance(idx, vp, chunk);
    const RmqVPlane v = rmq_vplane(vp, planes, H, W, bands);
    const RmqItem it = rmq_item(v.nc, chunk, chunks, channels, ws);
    if (vp != cur_vp) {
      if (cur_vp >= 0)
        rk_slow_pass<T, A, kMoi, kKeyed>(plane, tabs, ws, input, cur, r_lo, r_hi, ungrouped, channels, hw, H, W, PH, PW,
                                         bands, output, argmax, band_list ? chunks : 0);
      __syncthreads();  // everyone is done with the old plane and tables
      // ---- unpack: staged bytes (or global memory) -> fp32 plane, sanitised; min / max of the plane
      float lo = FLT_MAX, hi = -FLT_MAX;
      if (in_flight == vp) {
        mbar_wait(bar, parity);
        parity ^= 1u;
        unsigned src_off, off, bytes;
        window(vp, src_off, off, bytes);
        const T* sp = (const T*)(stage + off);
        for (int p = threadIdx.x; p < v.cells; p += kRmqThreads) {
          float x = Elem<T>::to_f(sp[p]);
          x = x > -FLT_MAX ? x : -FLT_MAX;
          plane[p] = x;
          lo = fminf(lo, x), hi = fmaxf(hi, x);
        }
      } else {
        const T* src = input + v.elem_off;
        for (int p = threadIdx.x; p < v.cells; p += kRmqThreads) {
          float x = Elem<T>::to_f(src[p]);
          x = x > -FLT_MAX ? x : -FLT_MAX;
          plane[p] = x;
          lo = fminf(lo, x), hi = fmaxf(hi, x);
        }
      }
      if (kBanded)
        for (int p = v.cells + threadIdx.x; p < hw; p += kRmqThreads) plane[p] = -FLT_MAX;  // a shorter last band
      if (kKeyed) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
          hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if ((threadIdx.x & 31) == 0) red[(threadIdx.x >> 5) * 2] = lo, red[(threadIdx.x >> 5) * 2 + 1] = hi;
      }
      __syncthreads();  // plane complete, staging buffer free, partials visible
      if (kKeyed && threadIdx.x < 32) {
        lo = red[threadIdx.x * 2], hi = red[threadIdx.x * 2 + 1];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
          hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (threadIdx.x == 0) {
          const float d = __fsub_rn(hi, lo);
          stats[0] = lo;
          stats[1] = d > 0.f ? __fdiv_rn(1022.f, d) : 0.f;  // d = +inf (overflow, +inf cells) -> 0: every key 1
        }
      }
      // the next plane this CTA will serve streams in behind the build and the queries of this one
      in_flight = -1;
      if (staged) {
        const int nx = sch.next_plane(idx, vp, chunk);
        if (nx >= 0 && issue(nx)) in_flight = nx;
      }
      __syncthreads();
      cur_vp = vp;
      cur = v;
      r_lo = it.r0, r_hi = it.r0;
      if (band_list) r_lo = chunk, r_hi = chunk;  // chunk numbers served, for the slow pass
      if (kKeyed)
        rk_build_tables(plane, tabs, hw, W, stats[0], stats[1]);
      else
        rmq_build_tables(plane, tabs, hw, W);
    }
    if (it.nr > 0) r_hi = it.r0 + it.nr;
    if (band_list) {
      r_hi = chunk + 1;
      const int slot = it.n * kBandListBands + v.band;
      const int off = __ldg(ws.band_off + slot), cnt = __ldg(ws.band_off + slot + 1) - off;
      const int j0 = off + (int)((long long)cnt * chunk / chunks), j1 = off + (int)((long long)cnt * (chunk + 1) / chunks);
      rk_main_loop_list<T, A, kKeyed>(plane, tabs, ws, it.c, j0, j1, PP, cpp, (unsigned)W, (unsigned)hw, magic_pp,
                                      (unsigned)(v.row0 * W), output, argmax);
    } else if (!kMoi && !kKeyed && !kBanded && !ungrouped && rows_loop)
      rk_main_loop_rows<T, A>(plane, tabs, ws, it, PP, cpp, (unsigned)W, magic_pp, output, argmax);
    else if (ungrouped)
      rk_main_loop<T, A, true, kMoi, kKeyed, kBanded>(plane, tabs, ws, it, PP, cpp, (unsigned)W, (unsigned)hw, magic_pp,
                                             moi.mask_scale, moi.zero_floor, bands.nb, v.band, (unsigned)(v.row0 * W),
                                             output, argmax);
    else
      rk_main_loop<T, A, false, kMoi, kKeyed, kBanded>(plane, tabs, ws, it, PP, cpp, (unsigned)W, (unsigned)hw, magic_pp,
                                              moi.mask_scale, moi.zero_floor, bands.nb, v.band, (unsigned)(v.row0 * W),
                                              output, argmax);
  }
  if (cur_vp >= 0)
    rk_slow_pass<T, A, kMoi, kKeyed>(plane, tabs, ws, input, cur, r_lo, r_hi, ungrouped, channels, hw, H, W, PH, PW, bands,
                                     output, argmax, band_list ? chunks : 0);
}

}  // namespace jtsm
