// RoIAlign forward / backward for sm_100a.
//
// Reference: detectron2/layers/roi_align.py:22-59,63-113 (module + autograd fn),
// detectron2/layers/csrc/ROIAlign/ROIAlign_cuda.cu:12-63 (bilinear_interpolate), :66-139 (forward),
// :142-306 (backward), CPU twin ROIAlign_cpu.cpp:21-114,117-218,287-395.  torchvision's roi_align
// (what roi_align.py:5-15 aliases when torchvision >= 0.7) computes the same values bit for bit
// (tests/test_oracle_pool.py).
//
// Design: same plane-per-CTA scheme as RoIPool -- one persistent CTA per SM walks (plane, RoI-chunk)
// items with the fp32 plane staged in shared memory (read once from HBM/L2, coalesced), RoI geometry
// precomputed once per call (channel independent).  Arithmetic follows the reference's operation
// order with explicit round-to-nearest intrinsics (no FMA contraction), so the forward is bit-exact
// against the CPU oracle; the backward uses RED.ADD.F32 (sum order differs -> tolerance).
#include <cub/cub.cuh>

#include "common.cuh"

namespace jtsm {

struct AlignGeom {  // ROIAlign_cpu.cpp:139-166
  float start_h, start_w, bin_h, bin_w;
  int grid_h, grid_w, batch, pad;
};

struct AlignWorkspace {
  AlignGeom* geom;  // [R]   (geom.pad = 1 when the RoI has tap tables)
  int* first;       // [batch]
  int* neg_last;    // [batch]
  int* ungrouped;   // [1]
  float* taps;      // [R][tap_stride]: PH * grid_h row sample coordinates, then PW * grid_w column ones (forward)
};

// Largest sampling grid per axis that gets a table, and the table entries of one RoI.  A RoI that lies inside the
// map has grid <= ceil(L / P) with the adaptive grid; bigger grids (boxes far outside the map, large fixed
// sampling ratios on tiny maps) take the direct path.  P * cap <= L + 2 * P, which bounds the workspace without
// knowing P (jtsm_roi_align_workspace_bytes has no pooled-size arguments; P <= 64).
__host__ __device__ inline int align_grid_cap(int L, int P) { return L / P + 2; }
__host__ __device__ inline int align_tap_stride(int H, int W, int PH, int PW) {
  return PH * align_grid_cap(H, PH) + PW * align_grid_cap(W, PW);
}
inline size_t align_tap_bytes(int H, int W, int num_rois) {
  return align_up((size_t)num_rois * ((size_t)H + W + 4 * 64) * sizeof(float), 256);
}

inline size_t align_workspace_bytes(int batch, int num_rois) {
  return align_up((size_t)num_rois * sizeof(AlignGeom), 256) + align_up((size_t)batch * 8 + 16, 256) + 256;
}
inline AlignWorkspace carve_align_workspace(void* ws, int batch, int num_rois) {
  char* p = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  AlignWorkspace w;
  w.geom = (AlignGeom*)p;
  p += align_up((size_t)num_rois * sizeof(AlignGeom), 256);
  w.first = (int*)p;
  w.neg_last = w.first + batch;
  w.ungrouped = w.neg_last + batch;
  p += align_up((size_t)batch * 8 + 16, 256);
  w.taps = (float*)p;  // only valid when the caller sized the workspace for the forward pass
  return w;
}

__device__ __forceinline__ AlignGeom align_geom(const float* roi, float scale, int PH, int PW,
                                                int sampling_ratio, int aligned) {
  AlignGeom g;
  g.batch = (int)roi[0];
  const float off = aligned ? 0.5f : 0.f;
  const float sw = __fsub_rn(__fmul_rn(roi[1], scale), off), sh = __fsub_rn(__fmul_rn(roi[2], scale), off);
  const float ew = __fsub_rn(__fmul_rn(roi[3], scale), off), eh = __fsub_rn(__fmul_rn(roi[4], scale), off);
  float rw = __fsub_rn(ew, sw), rh = __fsub_rn(eh, sh);
  if (!aligned) {
    rw = fmaxf(rw, 1.f);
    rh = fmaxf(rh, 1.f);
  }
  g.start_h = sh;
  g.start_w = sw;
  g.bin_h = __fdiv_rn(rh, (float)PH);
  g.bin_w = __fdiv_rn(rw, (float)PW);
  g.grid_h = sampling_ratio > 0 ? sampling_ratio : (int)ceilf(__fdiv_rn(rh, (float)PH));
  g.grid_w = sampling_ratio > 0 ? sampling_ratio : (int)ceilf(__fdiv_rn(rw, (float)PW));
  g.pad = 0;
  return g;
}

__global__ void roi_align_prep_kernel(const float* __restrict__ rois, int num_rois, int batch, float scale,
                                      int PH, int PW, int sampling_ratio, int aligned, AlignWorkspace ws) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= num_rois) return;
  AlignGeom g = align_geom(rois + (size_t)r * 5, scale, PH, PW, sampling_ratio, aligned);
  const bool ok = g.batch >= 0 && g.batch < batch;
  if (!ok) g.batch = -1;
  ws.geom[r] = g;
  if (r > 0 && (int)rois[(size_t)(r - 1) * 5] > g.batch) *ws.ungrouped = 1;
  if (ok) {
    atomicMin(ws.first + g.batch, r);
    atomicMin(ws.neg_last + g.batch, -r);
  }
}

// One bilinear sample: 4 taps + weights (ROIAlign_cuda.cu:12-63).  Returns false when the sample is
// outside the map (contributes 0).
struct Taps {
  int p1, p2, p3, p4;
  float w1, w2, w3, w4;
};
__device__ __forceinline__ bool bilinear_taps(int H, int W, float y, float x, Taps& t) {
  if (y < -1.f || y > (float)H || x < -1.f || x > (float)W) return false;
  if (y <= 0.f) y = 0.f;
  if (x <= 0.f) x = 0.f;
  int yl = (int)y, xl = (int)x, yh, xh;
  if (yl >= H - 1) {
    yh = yl = H - 1;
    y = (float)yl;
  } else {
    yh = yl + 1;
  }
  if (xl >= W - 1) {
    xh = xl = W - 1;
    x = (float)xl;
  } else {
    xh = xl + 1;
  }
  const float ly = __fsub_rn(y, (float)yl), lx = __fsub_rn(x, (float)xl);
  const float hy = __fsub_rn(1.f, ly), hx = __fsub_rn(1.f, lx);
  t.w1 = __fmul_rn(hy, hx);
  t.w2 = __fmul_rn(hy, lx);
  t.w3 = __fmul_rn(ly, hx);
  t.w4 = __fmul_rn(ly, lx);
  t.p1 = yl * W + xl;
  t.p2 = yl * W + xh;
  t.p3 = yh * W + xl;
  t.p4 = yh * W + xh;
  return true;
}

__device__ __forceinline__ float sample_coord(float start, int p, float bin, int i, int grid) {
  // start + p * bin + (i + .5f) * bin / grid      (ROIAlign_cuda.cu:119-126)
  return __fadd_rn(__fadd_rn(start, __fmul_rn((float)p, bin)),
                   __fdiv_rn(__fmul_rn(__fadd_rn((float)i, 0.5f), bin), (float)grid));
}

// Sample coordinate tables.  Everything about a sample that needs a division -- start + p * bin + (i + .5) * bin /
// grid -- is channel independent, so the forward kernel reads it from a per-RoI table instead of recomputing it for
// each of the C planes: one float per sample position and axis, already clamped at 0, kTapOutside when the sample is
// off the map.  (A 16-byte entry with the two cell offsets and both weights was tried first: the LDG.128 per sample
// costs 4 LSU wavefronts and the kernel is LSU bound -- profiles/r1q_roi_align.md; 4-byte entries cost one.)
constexpr float kTapOutside = -2.f;
__device__ __forceinline__ float align_tap_coord(float c, int L) {
  if (c < -1.f || c > (float)L) return kTapOutside;
  return c <= 0.f ? 0.f : c;
}
// cells and weights of a table coordinate: the tail of bilinear_taps for one axis
__device__ __forceinline__ void align_tap_cells(float c, int L, int& lo, int& hi, float& h, float& l) {
  lo = (int)c;
  if (lo >= L - 1) {
    hi = lo = L - 1;
    c = (float)lo;
  } else {
    hi = lo + 1;
  }
  l = __fsub_rn(c, (float)lo);
  h = __fsub_rn(1.f, l);
}

// Tables of every RoI whose grids fit (one CTA per RoI); marks the RoI in geom.pad.
__global__ void roi_align_taps_kernel(int num_rois, int H, int W, int PH, int PW, AlignWorkspace ws) {
  const int r = blockIdx.x;
  const AlignGeom g = ws.geom[r];
  if (g.batch < 0 || g.grid_h <= 0 || g.grid_w <= 0 || g.grid_h > align_grid_cap(H, PH) ||
      g.grid_w > align_grid_cap(W, PW))
    return;
  float* t = ws.taps + (size_t)r * align_tap_stride(H, W, PH, PW);
  const int ny = PH * g.grid_h, nx = PW * g.grid_w;
  for (int e = threadIdx.x; e < ny + nx; e += blockDim.x) {
    if (e < ny) {
      const int p = e / g.grid_h, i = e - p * g.grid_h;
      t[e] = align_tap_coord(sample_coord(g.start_h, p, g.bin_h, i, g.grid_h), H);
    } else {
      const int p = (e - ny) / g.grid_w, i = (e - ny) - p * g.grid_w;
      t[e] = align_tap_coord(sample_coord(g.start_w, p, g.bin_w, i, g.grid_w), W);
    }
  }
  if (threadIdx.x == 0) ws.geom[r].pad = 1;
}

constexpr int kAlignThreads = 512;

// grid: persistent CTAs over (plane, chunk) items.  kSmem: plane staged in shared memory.
// kPairs: the staged plane holds (p[y][x], p[y][min(x + 1, W - 1)]) pairs, so the two taps of a row come from ONE 8-byte
// shared load (the reference's x_high is exactly min(x_low + 1, W - 1)): 2 loads per sample instead of 4.
template <typename T, bool kSmem, bool kPairs = false>
__global__ void __launch_bounds__(kAlignThreads)
roi_align_fwd_kernel(const T* __restrict__ input, int channels, int H, int W, int PH, int PW, int chunks,
                     int num_items, AlignWorkspace ws, T* __restrict__ output) {
  extern __shared__ float plane_s[];
  const int hw = H * W;
  const int PP = PH * PW;
  const unsigned magic_pw = (unsigned)((0x100000000ull + PW - 1) / PW);
  const int step_r = kAlignThreads / PP, step_b = kAlignThreads - step_r * PP;
  const int rl_init = threadIdx.x / PP, bin_init = threadIdx.x - rl_init * PP;
  const bool ungrouped = *ws.ungrouped != 0;
  const int tap_stride = align_tap_stride(H, W, PH, PW);
  const int per = num_items / gridDim.x, rem = num_items % gridDim.x;
  const int i0 = blockIdx.x * per + min((int)blockIdx.x, rem);
  const int i1 = i0 + per + ((int)blockIdx.x < rem ? 1 : 0);
  int cur_plane = -1;
  for (int item = i0; item < i1; ++item) {
    const int nc = item / chunks;
    const int chunk = item - nc * chunks;
    const int n = nc / channels;
    const int c = nc - n * channels;
    const T* src = input + (size_t)nc * hw;
    if (kSmem && nc != cur_plane) {
      __syncthreads();
      cur_plane = nc;
      if (kPairs) {
        float2* pp = (float2*)plane_s;
        for (int p = threadIdx.x; p < hw; p += kAlignThreads) {
          const int x = p % W;
          pp[p] = make_float2(Elem<T>::to_f(src[p]), Elem<T>::to_f(src[x + 1 < W ? p + 1 : p]));
        }
      } else {
        for (int p = threadIdx.x; p < hw; p += kAlignThreads) plane_s[p] = Elem<T>::to_f(src[p]);
      }
      __syncthreads();
    }
    const int first = ws.first[n];
    const int last = -ws.neg_last[n];
    if (first > last) continue;
    const int len = last - first + 1;
    const int r0 = first + (int)(((long long)len * chunk) / chunks);
    const int nr = first + (int)(((long long)len * (chunk + 1)) / chunks) - r0;
    const size_t cpp = (size_t)channels * PP;
    int rl = rl_init, bin = bin_init;
    while (rl < nr) {
      const int r = r0 + rl;
      const int ph = (int)__umulhi((unsigned)bin, magic_pw);
      const int pw = bin - ph * PW;
      const size_t o = (size_t)r * cpp + (size_t)c * PP + bin;
      bin += step_b;
      rl += step_r;
      if (bin >= PP) {
        bin -= PP;
        ++rl;
      }
      const AlignGeom g = ws.geom[r];
      if (ungrouped && g.batch != n) continue;
      const float count = (float)max(g.grid_h * g.grid_w, 1);
      float v = 0.f;
      if (g.pad) {  // coordinate tables: no divisions left; products and sums in the reference's order
        const float* ty = ws.taps + (size_t)r * tap_stride + ph * g.grid_h;
        const float* tx = ws.taps + (size_t)r * tap_stride + PH * g.grid_h + pw * g.grid_w;
        for (int iy = 0; iy < g.grid_h; ++iy) {
          const float y = __ldg(ty + iy);
          if (y < 0.f) continue;  // adding the 0 of an outside sample changes nothing (v is never -0)
          int yl, yh;
          float hy, ly;
          align_tap_cells(y, H, yl, yh, hy, ly);
          yl *= W, yh *= W;
          for (int ix = 0; ix < g.grid_w; ++ix) {
            const float x = __ldg(tx + ix);
            if (x < 0.f) continue;
            int xl, xh;
            float hx, lx;
            align_tap_cells(x, W, xl, xh, hx, lx);
            float v1, v2, v3, v4;
            if (kPairs) {
              const float2 a = ((const float2*)plane_s)[yl + xl], b = ((const float2*)plane_s)[yh + xl];
              v1 = a.x; v2 = a.y; v3 = b.x; v4 = b.y;
            } else if (kSmem) {
              v1 = plane_s[yl + xl]; v2 = plane_s[yl + xh]; v3 = plane_s[yh + xl]; v4 = plane_s[yh + xh];
            } else {
              v1 = Elem<T>::to_f(__ldg(src + yl + xl)); v2 = Elem<T>::to_f(__ldg(src + yl + xh));
              v3 = Elem<T>::to_f(__ldg(src + yh + xl)); v4 = Elem<T>::to_f(__ldg(src + yh + xh));
            }
            const float w1 = __fmul_rn(hy, hx), w2 = __fmul_rn(hy, lx);
            const float w3 = __fmul_rn(ly, hx), w4 = __fmul_rn(ly, lx);
            v = __fadd_rn(v, __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w1, v1), __fmul_rn(w2, v2)), __fmul_rn(w3, v3)),
                                       __fmul_rn(w4, v4)));
          }
        }
        output[o] = Elem<T>::from_f(__fdiv_rn(v, count));
        continue;
      }
      for (int iy = 0; iy < g.grid_h; ++iy) {
        const float y = sample_coord(g.start_h, ph, g.bin_h, iy, g.grid_h);
        for (int ix = 0; ix < g.grid_w; ++ix) {
          const float x = sample_coord(g.start_w, pw, g.bin_w, ix, g.grid_w);
          Taps t;
          float val = 0.f;
          if (bilinear_taps(H, W, y, x, t)) {
            float v1, v2, v3, v4;
            if (kPairs) {
              const float2 a = ((const float2*)plane_s)[t.p1], b = ((const float2*)plane_s)[t.p3];
              v1 = a.x; v2 = a.y; v3 = b.x; v4 = b.y;
            } else if (kSmem) {
              v1 = plane_s[t.p1]; v2 = plane_s[t.p2]; v3 = plane_s[t.p3]; v4 = plane_s[t.p4];
            } else {
              v1 = Elem<T>::to_f(__ldg(src + t.p1)); v2 = Elem<T>::to_f(__ldg(src + t.p2));
              v3 = Elem<T>::to_f(__ldg(src + t.p3)); v4 = Elem<T>::to_f(__ldg(src + t.p4));
            }
            val = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(t.w1, v1), __fmul_rn(t.w2, v2)), __fmul_rn(t.w3, v3)),
                            __fmul_rn(t.w4, v4));
          }
          v = __fadd_rn(v, val);
        }
      }
      output[o] = Elem<T>::from_f(__fdiv_rn(v, count));
    }
  }
}

// backward: one thread per output element (bin fastest), 4 RED.ADD.F32 per sample
template <typename T>
__global__ void __launch_bounds__(256)
roi_align_bwd_kernel(const T* __restrict__ grad_output, long long sn, long long sc, long long sh, long long sw,
                     int channels, int H, int W, int PH, int PW, int num_rois, AlignWorkspace ws,
                     float* __restrict__ grad_acc) {
  const int PP = PH * PW;
  const int hw = H * W;
  const unsigned magic_pw = (unsigned)((0x100000000ull + PW - 1) / PW);
  const unsigned magic_pp = (unsigned)((0x100000000ull + PP - 1) / PP);
  const int r = blockIdx.x;
  const AlignGeom g = ws.geom[r];
  if (g.batch < 0) return;
  const float count = (float)(g.grid_h * g.grid_w);  // unclamped in backward (ROIAlign_cuda.cu:258)
  const int cpp = channels * PP;
  float* acc_b = grad_acc + (size_t)g.batch * channels * hw;
  for (int i = blockIdx.y * 256 + threadIdx.x; i < cpp; i += gridDim.y * 256) {
    const int c = (int)__umulhi((unsigned)i, magic_pp);
    const int bin = i - c * PP;
    const int ph = (int)__umulhi((unsigned)bin, magic_pw);
    const int pw = bin - ph * PW;
    const float gb = Elem<T>::to_f(grad_output[r * sn + c * sc + ph * sh + pw * sw]);
    float* acc = acc_b + (size_t)c * hw;
    for (int iy = 0; iy < g.grid_h; ++iy) {
      const float y = sample_coord(g.start_h, ph, g.bin_h, iy, g.grid_h);
      for (int ix = 0; ix < g.grid_w; ++ix) {
        const float x = sample_coord(g.start_w, pw, g.bin_w, ix, g.grid_w);
        Taps t;
        if (!bilinear_taps(H, W, y, x, t)) continue;
        atomicAdd(acc + t.p1, __fdiv_rn(__fmul_rn(gb, t.w1), count));
        atomicAdd(acc + t.p2, __fdiv_rn(__fmul_rn(gb, t.w2), count));
        atomicAdd(acc + t.p3, __fdiv_rn(__fmul_rn(gb, t.w3), count));
        atomicAdd(acc + t.p4, __fdiv_rn(__fmul_rn(gb, t.w4), count));
      }
    }
  }
}

// backward, separable form: the bilinear weight of a sample is (row weight) x (column weight), and both depend on
// one axis only, so   grad_in[y][x] += 1/count * sum_ph sum_pw  Wy[y][ph] * g[ph][pw] * Wx[x][pw]
// with Wy[y][ph] = sum of the row weights the samples of bin row ph put on map row y (ROIAlign_cuda.cu:203-306
// distributes g * w_i / count per sample; same sum, different association -> tolerance like any atomic backward).
// One CTA per (RoI, channel slice): the two weight matrices are built once (channel independent), the touched
// rows / columns are compacted, and every touched cell of every channel gets ONE RED instead of one per tap
// (16 per bin at sampling_ratio 2, 4 * grid^2 per bin with the adaptive grid).
// dynamic smem: Wy [H][PH] | Wx [W][PW] | row list [H] | column list [W]   (list entry: y | p_lo << 16 | p_hi << 24)
constexpr int kAlignBwdThreads = 256;

__device__ __forceinline__ void align_axis_weights(float start, float bin, int grid, int P, int L, float* w,
                                                   int stride, int* list, int* count_out) {
  // called by one full warp; lane p owns column p of w, so its sequential adds need no atomics
  const int lane = threadIdx.x & 31;
  int lo_min = L, hi_max = -1;
  for (int p = lane; p < P; p += 32) {
    for (int i = 0; i < grid; ++i) {
      float c = sample_coord(start, p, bin, i, grid);
      if (c < -1.f || c > (float)L) continue;
      if (c <= 0.f) c = 0.f;
      int lo = (int)c, hi;
      if (lo >= L - 1) {
        hi = lo = L - 1;
        c = (float)lo;
      } else {
        hi = lo + 1;
      }
      const float l = __fsub_rn(c, (float)lo), h = __fsub_rn(1.f, l);
      w[lo * stride + p] += h;
      w[hi * stride + p] += l;
      lo_min = min(lo_min, lo);
      hi_max = max(hi_max, hi);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo_min = min(lo_min, __shfl_xor_sync(0xffffffffu, lo_min, o));
    hi_max = max(hi_max, __shfl_xor_sync(0xffffffffu, hi_max, o));
  }
  __syncwarp();
  int n = 0;
  for (int base = lo_min; base <= hi_max; base += 32) {
    const int y = base + lane;
    int p_lo = P, p_hi = -1;
    if (y <= hi_max)
      for (int p = 0; p < P; ++p)
        if (w[y * stride + p] != 0.f) {
          p_lo = min(p_lo, p);
          p_hi = p;
        }
    const unsigned m = __ballot_sync(0xffffffffu, p_hi >= 0);
    if (p_hi >= 0) list[n + __popc(m & ((1u << lane) - 1u))] = y | (p_lo << 16) | (p_hi << 24);
    n += __popc(m);
  }
  if (lane == 0) *count_out = n;
}

template <typename T>
__global__ void __launch_bounds__(kAlignBwdThreads)
roi_align_bwd_patch_kernel(const T* __restrict__ grad_output, long long sn, long long sc, long long sh, long long sw,
                           int channels, int H, int W, int PH, int PW, AlignWorkspace ws,
                           float* __restrict__ grad_acc) {
  extern __shared__ float align_bwd_s[];
  float* wy = align_bwd_s;
  float* wx = wy + H * PH;
  int* rows = (int*)(wx + W * PW);
  int* cols = rows + H;
  __shared__ int s_n[2];
  const int r = blockIdx.x;
  const AlignGeom g = ws.geom[r];
  if (g.batch < 0 || g.grid_h <= 0 || g.grid_w <= 0) return;
  for (int i = threadIdx.x; i < H * PH + W * PW; i += kAlignBwdThreads) wy[i] = 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) align_axis_weights(g.start_h, g.bin_h, g.grid_h, PH, H, wy, PH, rows, &s_n[0]);
  if (warp == 1) align_axis_weights(g.start_w, g.bin_w, g.grid_w, PW, W, wx, PW, cols, &s_n[1]);
  __syncthreads();
  const int nr = s_n[0], nc = s_n[1], cells = nr * nc;
  if (cells == 0) return;
  // channel slice of this CTA
  const int c0 = (int)(((long long)channels * blockIdx.y) / gridDim.y);
  const int c1 = (int)(((long long)channels * (blockIdx.y + 1)) / gridDim.y);
  const float inv = __fdiv_rn(1.f, (float)(g.grid_h * g.grid_w));  // unclamped count (ROIAlign_cuda.cu:258)
  // item = (channel, row entry, column entry), column fastest; stepped by the CTA size without divisions
  int cc = c0 + threadIdx.x / cells;
  int rem = threadIdx.x % cells;
  int jj = rem / nc, kk = rem - jj * nc;
  const int step_c = kAlignBwdThreads / cells, srem = kAlignBwdThreads % cells;
  const int step_j = srem / nc, step_k = srem - step_j * nc;
  const T* gr = grad_output + r * sn;
  float* acc_b = grad_acc + (size_t)g.batch * channels * H * W;
  while (cc < c1) {
    const int ri = rows[jj], ci = cols[kk];
    const int y = ri & 0xFFFF, p_lo = (ri >> 16) & 0xFF, p_hi = (ri >> 24) & 0xFF;
    const int x = ci & 0xFFFF, q_lo = (ci >> 16) & 0xFF, q_hi = (ci >> 24) & 0xFF;
    const T* gp = gr + cc * sc;
    const float* wyr = wy + y * PH;
    const float* wxr = wx + x * PW;
    float sum = 0.f;
    for (int ph = p_lo; ph <= p_hi; ++ph) {
      const T* gq = gp + ph * sh;
      float inner = 0.f;
      for (int pw = q_lo; pw <= q_hi; ++pw) inner = fmaf(Elem<T>::to_f(gq[pw * sw]), wxr[pw], inner);
      sum = fmaf(wyr[ph], inner, sum);
    }
    if (sum != 0.f) atomicAdd(acc_b + ((size_t)cc * H + y) * W + x, sum * inv);
    kk += step_k;
    if (kk >= nc) {
      kk -= nc;
      ++jj;
    }
    jj += step_j;
    if (jj >= nr) {
      jj -= nr;
      ++cc;
    }
    cc += step_c;
  }
}

// Same, for PH <= PMAX (8 or 16): a thread owns one touched column x of one channel, folds the column weights into
// PH partial sums   t[ph] = sum_pw g[ph][pw] * Wx[x][pw]   once (registers), then walks the touched rows with one
// broadcast read of the row's weights, PMAX FMAs and one RED per cell; neighbouring threads own neighbouring
// columns, so the REDs of a warp fall into a few sectors.  (Weights outside a row's bins are exact zeros; a
// non-finite gradient therefore spreads over the RoI's column instead of its own bins only.)
// dynamic smem: Wy [H][PMAX] | Wx [W][PW] | row list [H] | column list [W]
template <typename T, int PMAX>
__global__ void __launch_bounds__(kAlignBwdThreads)
roi_align_bwd_cols_kernel(const T* __restrict__ grad_output, long long sn, long long sc, long long sh, long long sw,
                          int channels, int H, int W, int PH, int PW, AlignWorkspace ws,
                          float* __restrict__ grad_acc) {
  extern __shared__ __align__(16) float align_bwd_s[];
  float* wy = align_bwd_s;
  float* wx = wy + H * PMAX;
  int* rows = (int*)(wx + W * PW);
  int* cols = rows + H;
  __shared__ int s_n[2];
  const int r = blockIdx.x;
  const AlignGeom g = ws.geom[r];
  if (g.batch < 0 || g.grid_h <= 0 || g.grid_w <= 0) return;
  for (int i = threadIdx.x; i < H * PMAX + W * PW; i += kAlignBwdThreads) wy[i] = 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) align_axis_weights(g.start_h, g.bin_h, g.grid_h, PH, H, wy, PMAX, rows, &s_n[0]);
  if (warp == 1) align_axis_weights(g.start_w, g.bin_w, g.grid_w, PW, W, wx, PW, cols, &s_n[1]);
  __syncthreads();
  const int nr = s_n[0], nc = s_n[1];
  if (nr == 0 || nc == 0) return;
  const int c0 = (int)(((long long)channels * blockIdx.y) / gridDim.y);
  const int c1 = (int)(((long long)channels * (blockIdx.y + 1)) / gridDim.y);
  const float inv = __fdiv_rn(1.f, (float)(g.grid_h * g.grid_w));  // unclamped count (ROIAlign_cuda.cu:258)
  int cc = c0 + threadIdx.x / nc, kk = threadIdx.x % nc;
  const int step_c = kAlignBwdThreads / nc, step_k = kAlignBwdThreads % nc;
  const T* gr = grad_output + r * sn;
  float* acc_b = grad_acc + (size_t)g.batch * channels * H * W;
  while (cc < c1) {
    const int ci = cols[kk];
    const int x = ci & 0xFFFF, q_lo = (ci >> 16) & 0xFF, q_hi = (ci >> 24) & 0xFF;
    const T* gp = gr + cc * sc;
    const float* wxr = wx + x * PW;
    float t[PMAX];
#pragma unroll
    for (int ph = 0; ph < PMAX; ++ph) t[ph] = 0.f;
    for (int pw = q_lo; pw <= q_hi; ++pw) {
      const float wv = wxr[pw];
      const T* gq = gp + pw * sw;
#pragma unroll
      for (int ph = 0; ph < PMAX; ++ph)
        if (ph < PH) t[ph] = fmaf(Elem<T>::to_f(gq[ph * sh]), wv, t[ph]);
    }
    float* acc = acc_b + (size_t)cc * H * W + x;
    for (int jj = 0; jj < nr; ++jj) {
      const int y = rows[jj] & 0xFFFF;
      const float4* wr = reinterpret_cast<const float4*>(wy + y * PMAX);
      float sum = 0.f;
#pragma unroll
      for (int q = 0; q < PMAX / 4; ++q) {
        const float4 w4 = wr[q];
        sum = fmaf(w4.x, t[4 * q], sum);
        sum = fmaf(w4.y, t[4 * q + 1], sum);
        sum = fmaf(w4.z, t[4 * q + 2], sum);
        sum = fmaf(w4.w, t[4 * q + 3], sum);
      }
      if (sum != 0.f) atomicAdd(acc + y * W, sum * inv);
    }
    kk += step_k;
    if (kk >= nc) {
      kk -= nc;
      ++cc;
    }
    cc += step_c;
  }
}

template <typename T>
__global__ void align_convert_from_f32_kernel(const float* __restrict__ src, T* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = Elem<T>::from_f(src[i]);
}

}  // namespace jtsm

using namespace jtsm;

namespace {
int align_prep(const float* rois, int num_rois, int batch, float scale, int PH, int PW, int sampling_ratio,
               int aligned, const AlignWorkspace& ws, cudaStream_t st, int H = 0, int W = 0, bool taps = false) {
  cudaMemsetAsync(ws.first, 0x7f, (size_t)batch * 2 * sizeof(int), st);
  cudaMemsetAsync(ws.ungrouped, 0, sizeof(int), st);
  roi_align_prep_kernel<<<(num_rois + 127) / 128, 128, 0, st>>>(rois, num_rois, batch, scale, PH, PW,
                                                                sampling_ratio, aligned, ws);
  count_launch();
  if (taps) {
    roi_align_taps_kernel<<<num_rois, 64, 0, st>>>(num_rois, H, W, PH, PW, ws);
    count_launch();
  }
  return cuda_status();
}

template <typename T>
int launch_align_fwd(const void* input, int batch, int channels, int H, int W, int num_rois, int PH, int PW,
                     const AlignWorkspace& ws, void* output, cudaStream_t st) {
  const int planes = batch * channels;
  const size_t plane_b = (size_t)H * W * sizeof(float);
  const bool use_smem = plane_b <= 200 * 1024 && !tuning("roi_align_force_gmem", 0);
  // pairs only while four CTAs still fit per SM: at c2 (30 KB planes) the lost occupancy costs more than the halved
  // load count saves (sr 2: 673 vs 614 us; adaptive 1972 vs 2085), at c1 (7.6 KB) it wins 5-7 % (profiles/r1q)
  const bool pairs = use_smem && 2 * plane_b <= 52 * 1024 && tuning("roi_align_fwd_pairs", 1) != 0;
  const size_t smem = pairs ? 2 * plane_b : plane_b;
  int per_sm = use_smem ? (int)((220 * 1024) / (smem + 1024)) : 4;
  per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  const int max_grid = num_sms() * per_sm;
  int chunks = (8 * max_grid + planes - 1) / planes;
  const int cap = num_rois / 32 > 1 ? num_rois / 32 : 1;
  chunks = chunks > cap ? cap : (chunks < 1 ? 1 : chunks);
  const int items = planes * chunks;
  const int grid = items < max_grid ? items : max_grid;
  if (use_smem) {
    auto k = pairs ? roi_align_fwd_kernel<T, true, true> : roi_align_fwd_kernel<T, true, false>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<grid, kAlignThreads, smem, st>>>((const T*)input, channels, H, W, PH, PW, chunks, items, ws, (T*)output);
  } else {
    roi_align_fwd_kernel<T, false><<<grid, kAlignThreads, 0, st>>>((const T*)input, channels, H, W, PH, PW,
                                                                    chunks, items, ws, (T*)output);
  }
  count_launch();
  return cuda_status();
}
}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// Deterministic backward: a segmented reduction instead of atomics (north_star; fixed sampling_ratio).
// The tap pattern of RoIAlign is channel independent, so it is resolved once per call:
//   1. `align_det_taps_kernel`: every (RoI, bin, sample, corner) tap -> key = image * H*W + cell, payload = (output
//      element r * PP + bin, bilinear weight); tap id = ((((r * PH + ph) * PW + pw) * gh + iy) * gw + ix) * 4 + corner;
//   2. a STABLE radix sort by key (CUB): the taps of a cell end up contiguous and in tap-id order -- exactly the order
//      in which the reference's serial CPU backward (ROIAlign_cpu.cpp:168-263: for n, c, ph, pw, iy, ix; add w1..w4)
//      adds to that cell;  `align_det_bounds_kernel`: first tap of every cell by binary search;
//   3. `align_det_transpose_kernel`: grad_output (any strides, any dtype) -> fp32 [R * PP][C], channel last;
//   4. `align_det_reduce_kernel`: a warp per cell, lanes along channels: acc += (g * w) / count tap by tap in list
//      order (coalesced 128-byte reads of the transposed gradient), then a shared-memory transpose so the map is
//      written with lanes along cells.  No atomics, every element of grad_input written exactly once (no memset),
//      run-to-run bit-identical, and bit-EQUAL to the reference's CPU backward (same operands, same order).
// Cost: the sort and the 4-byte-per-channel tap walk make it ~2-3x slower than the RED kernels; it is an opt-in.
struct AlignDetTap {
  unsigned out;  // r * PP + bin
  float w;
};

__global__ void __launch_bounds__(256)
align_det_taps_kernel(int num_rois, int H, int W, int PH, int PW, int S, AlignWorkspace ws, unsigned sentinel,
                      unsigned* __restrict__ keys, unsigned long long* __restrict__ vals, float* __restrict__ counts) {
  const int PP = PH * PW, per_roi = PP * S * S;
  const long long total = (long long)num_rois * per_roi;
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const int r = (int)(i / per_roi);
  int rem = (int)(i - (long long)r * per_roi);
  const int bin = rem / (S * S);
  rem -= bin * S * S;
  const int iy = rem / S, ix = rem - iy * S;
  const int ph = bin / PW, pw = bin - ph * PW;
  const AlignGeom g = ws.geom[r];
  if (rem == 0 && bin == 0) counts[r] = (float)(g.grid_h * g.grid_w);
  Taps t;
  bool ok = g.batch >= 0;
  if (ok) {
    const float y = sample_coord(g.start_h, ph, g.bin_h, iy, g.grid_h);
    const float x = sample_coord(g.start_w, pw, g.bin_w, ix, g.grid_w);
    ok = bilinear_taps(H, W, y, x, t);
  }
  const unsigned base = ok ? (unsigned)g.batch * (unsigned)(H * W) : 0u;
  const unsigned out = (unsigned)r * (unsigned)PP + (unsigned)bin;
  const int pos[4] = {t.p1, t.p2, t.p3, t.p4};
  const float wt[4] = {t.w1, t.w2, t.w3, t.w4};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    keys[i * 4 + k] = ok ? base + (unsigned)pos[k] : sentinel;  // samples outside the map contribute nothing
    vals[i * 4 + k] = ((unsigned long long)out << 32) | (unsigned long long)__float_as_uint(ok ? wt[k] : 0.f);
  }
}

// first[k] = index of the first sorted tap with key >= k, for k in [0, cells]
__global__ void align_det_bounds_kernel(const unsigned* __restrict__ keys, long long num_taps, int cells,
                                        int* __restrict__ first) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > cells) return;
  long long lo = 0, hi = num_taps;
  while (lo < hi) {
    const long long mid = (lo + hi) >> 1;
    if (keys[mid] < (unsigned)k) lo = mid + 1; else hi = mid;
  }
  first[k] = (int)lo;
}

// gT[(r * PP + bin) * C + c] = float(grad_output[r, c, ph, pw]); 32 x 32 tiles through shared memory
template <typename T>
__global__ void __launch_bounds__(1024)
align_det_transpose_kernel(const T* __restrict__ g, long long sn, long long sc, long long sh, long long sw, int C,
                           int PH, int PW, float* __restrict__ gT) {
  __shared__ float tile[32][33];
  const int PP = PH * PW;
  const int r = blockIdx.z;
  const int c0 = blockIdx.y * 32, b0 = blockIdx.x * 32;
  {
    const int c = c0 + threadIdx.y, bin = b0 + threadIdx.x;
    if (c < C && bin < PP) {
      const int ph = bin / PW, pw = bin - ph * PW;
      tile[threadIdx.y][threadIdx.x] = Elem<T>::to_f(g[r * sn + c * sc + ph * sh + pw * sw]);
    }
  }
  __syncthreads();
  const int bin = b0 + threadIdx.y, c = c0 + threadIdx.x;
  if (c < C && bin < PP) gT[((size_t)r * PP + bin) * C + c] = tile[threadIdx.x][threadIdx.y];
}

template <typename T>
__global__ void __launch_bounds__(1024)
align_det_reduce_kernel(const unsigned long long* __restrict__ vals, const int* __restrict__ first,
                        const float* __restrict__ counts, const float* __restrict__ gT, int C, int hw, int cells, int PP,
                        T* __restrict__ grad_input) {
  __shared__ float tile[32][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = blockIdx.x * 32 + warp;  // global cell id = image * hw + cell in image
  const int c = blockIdx.y * 32 + lane;
  const unsigned magic_pp = (unsigned)((0x100000000ull + PP - 1) / PP);
  float acc = 0.f;
  if (cell < cells) {
    const int t0 = first[cell], t1 = first[cell + 1];
    const bool live = c < C;
    for (int t = t0; t < t1; ++t) {
      const unsigned long long v = __ldg(vals + t);  // same address in every lane
      const unsigned out = (unsigned)(v >> 32);
      const float w = __uint_as_float((unsigned)v);
      const float cnt = __ldg(counts + __umulhi(out, magic_pp));
      const float gv = live ? __ldg(gT + (size_t)out * C + c) : 0.f;
      acc = __fadd_rn(acc, __fdiv_rn(__fmul_rn(gv, w), cnt));  // grad * w / count, ROIAlign_cpu.cpp:238-247
    }
  }
  tile[warp][lane] = acc;
  __syncthreads();
  // transposed write: lanes along cells
  const int oc = blockIdx.y * 32 + warp, ocell = blockIdx.x * 32 + lane;
  if (oc < C && ocell < cells) {
    const int n = ocell / hw, p = ocell - n * hw;
    grad_input[((size_t)n * C + oc) * hw + p] = Elem<T>::from_f(tile[lane][warp]);
  }
}

extern "C" size_t jtsm_roi_align_workspace_bytes(int dtype, int batch, int channels, int height, int width,
                                                 int num_rois, int backward) {
  if (batch < 0 || num_rois < 0) return 0;
  size_t b = align_workspace_bytes(batch, num_rois);
  if (!backward) b += align_tap_bytes(height, width, num_rois);
  if (backward && dtype != JTSM_F32) b += (size_t)batch * channels * height * width * sizeof(float) + 256;
  return b;
}

extern "C" int jtsm_roi_align_forward(const void* input, const float* rois, int dtype, int batch, int channels,
                                      int height, int width, int num_rois, float spatial_scale, int pooled_h,
                                      int pooled_w, int sampling_ratio, int aligned, void* output,
                                      void* workspace, size_t workspace_bytes, void* stream) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0 ||
      pooled_h > 64 || pooled_w > 64)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois == 0 || channels == 0) return JTSM_OK;
  if (!input || !rois || !output || !workspace || batch == 0 || height == 0 || width == 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if ((long long)height * width > (1 << 30)) return JTSM_ERR_UNSUPPORTED;
  if (workspace_bytes < jtsm_roi_align_workspace_bytes(dtype, batch, channels, height, width, num_rois, 0))
    return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  AlignWorkspace ws = carve_align_workspace(workspace, batch, num_rois);
  int rc = align_prep(rois, num_rois, batch, spatial_scale, pooled_h, pooled_w, sampling_ratio, aligned, ws, st,
                      height, width, tuning("roi_align_fwd_algo", 0) != 1);
  if (rc != JTSM_OK) return rc;
  JTSM_DISPATCH_DTYPE(dtype, T, {
    rc = launch_align_fwd<T>(input, batch, channels, height, width, num_rois, pooled_h, pooled_w, ws, output, st);
  });
  return rc;
}

extern "C" int jtsm_roi_align_backward(const void* grad_output, long long stride_n, long long stride_c,
                                       long long stride_h, long long stride_w, const float* rois, int dtype,
                                       int batch, int channels, int height, int width, int num_rois,
                                       float spatial_scale, int pooled_h, int pooled_w, int sampling_ratio,
                                       int aligned, void* grad_input, void* workspace, size_t workspace_bytes,
                                       void* stream) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  const size_t n_in = (size_t)batch * channels * height * width;
  if (n_in == 0) return JTSM_OK;
  if (!grad_input || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_roi_align_workspace_bytes(dtype, batch, channels, height, width, num_rois, 1))
    return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  AlignWorkspace ws = carve_align_workspace(workspace, batch, num_rois);
  float* acc = (float*)grad_input;
  if (dtype != JTSM_F32)
    acc = (float*)(((uintptr_t)workspace + align_workspace_bytes(batch, num_rois) + 255) & ~(uintptr_t)255);
  cudaMemsetAsync(acc, 0, n_in * sizeof(float), st);
  int rc = JTSM_OK;
  if (num_rois > 0 && channels > 0) {
    if (!grad_output || !rois) return JTSM_ERR_INVALID_ARGUMENT;
    rc = align_prep(rois, num_rois, batch, spatial_scale, pooled_h, pooled_w, sampling_ratio, aligned, ws, st);
    if (rc != JTSM_OK) return rc;
    const int algo = tuning("roi_align_bwd_algo", 0);
    const int pmax = pooled_h <= 8 ? 8 : 16;
    const bool cols = algo != 2 && pooled_h <= 16;
    const size_t smem = ((size_t)height * (cols ? pmax : pooled_h) + (size_t)width * pooled_w + height + width) * 4;
    const bool patch = algo != 1 && smem <= 96 * 1024 && height < 65536 && width < 65536 && pooled_h < 256 &&
                       pooled_w < 256;
    if (patch) {
      // every CTA rebuilds the RoI's weight matrices, so only as many channel slices as it takes to fill the GPU
      int ysplit = (16 * num_sms() + num_rois - 1) / num_rois;
      ysplit = ysplit < 1 ? 1 : (ysplit > channels ? channels : ysplit);
      dim3 grid(num_rois, ysplit);
      JTSM_DISPATCH_DTYPE(dtype, T, {
        auto k = !cols ? roi_align_bwd_patch_kernel<T>
                       : (pmax == 8 ? roi_align_bwd_cols_kernel<T, 8> : roi_align_bwd_cols_kernel<T, 16>);
        if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, kAlignBwdThreads, smem, st>>>((const T*)grad_output, stride_n, stride_c, stride_h, stride_w,
                                                channels, height, width, pooled_h, pooled_w, ws, acc);
      });
    } else {
      const long long pp_ll = (long long)pooled_h * pooled_w;
      if ((long long)channels * pp_ll * pp_ll >= 0xFFFFFFFFLL) return JTSM_ERR_UNSUPPORTED;  // exact magic division
      const int cpp = channels * pooled_h * pooled_w;
      int ysplit = (cpp + 255) / 256;
      const int want = (4 * num_sms() * 8 + num_rois - 1) / num_rois;
      if (ysplit > want) ysplit = want < 1 ? 1 : want;
      dim3 grid(num_rois, ysplit);
      JTSM_DISPATCH_DTYPE(dtype, T, {
        roi_align_bwd_kernel<T><<<grid, 256, 0, st>>>((const T*)grad_output, stride_n, stride_c, stride_h, stride_w,
                                                      channels, height, width, pooled_h, pooled_w, num_rois, ws, acc);
      });
    }
    count_launch();
    rc = cuda_status();
    if (rc != JTSM_OK) return rc;
  }
  if (dtype == JTSM_BF16) {
    align_convert_from_f32_kernel<__nv_bfloat16><<<num_sms() * 8, 256, 0, st>>>(acc, (__nv_bfloat16*)grad_input, n_in);
    count_launch();
  } else if (dtype == JTSM_F16) {
    align_convert_from_f32_kernel<__half><<<num_sms() * 8, 256, 0, st>>>(acc, (__half*)grad_input, n_in);
    count_launch();
  }
  return cuda_status();
}

// ---- deterministic backward: launcher ----
namespace {
struct DetLayout {
  size_t taps, keys, vals, first, counts, gt, cub, total;
};
DetLayout det_layout(int batch, int channels, int height, int width, int num_rois, int PH, int PW, int S) {
  DetLayout L;
  L.taps = (size_t)num_rois * PH * PW * S * S * 4;
  L.keys = align_up(L.taps * sizeof(unsigned), 256);
  L.vals = align_up(L.taps * sizeof(unsigned long long), 256);
  L.first = align_up(((size_t)batch * height * width + 1) * sizeof(int), 256);
  L.counts = align_up((size_t)num_rois * sizeof(float), 256);
  L.gt = align_up((size_t)num_rois * PH * PW * channels * sizeof(float), 256);
  size_t tmp = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const unsigned*)nullptr, (unsigned*)nullptr,
                                  (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                  (long long)(L.taps ? L.taps : 1));
  L.cub = align_up(tmp, 256);
  L.total = align_workspace_bytes(batch, num_rois) + 2 * L.keys + 2 * L.vals + L.first + L.counts + L.gt + L.cub + 512;
  return L;
}
}  // namespace

extern "C" size_t jtsm_roi_align_backward_det_workspace_bytes(int batch, int channels, int height, int width,
                                                              int num_rois, int pooled_h, int pooled_w,
                                                              int sampling_ratio) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0 ||
      sampling_ratio <= 0)
    return 0;
  return det_layout(batch, channels, height, width, num_rois, pooled_h, pooled_w, sampling_ratio).total;
}

extern "C" int jtsm_roi_align_backward_det(const void* grad_output, long long stride_n, long long stride_c,
                                           long long stride_h, long long stride_w, const float* rois, int dtype,
                                           int batch, int channels, int height, int width, int num_rois,
                                           float spatial_scale, int pooled_h, int pooled_w, int sampling_ratio,
                                           int aligned, void* grad_input, void* workspace, size_t workspace_bytes,
                                           void* stream) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (sampling_ratio <= 0) return JTSM_ERR_UNSUPPORTED;  // the adaptive grid has no a-priori tap count
  const size_t n_in = (size_t)batch * channels * height * width;
  if (n_in == 0) return JTSM_OK;
  if (!grad_input || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  const long long cells_ll = (long long)batch * height * width;
  const DetLayout L = det_layout(batch, channels, height, width, num_rois, pooled_h, pooled_w, sampling_ratio);
  if (cells_ll >= 0x7FFFFFFFLL || L.taps >= 0x7FFFFFFFULL || (long long)num_rois * pooled_h * pooled_w >= 0x7FFFFFFFLL ||
      (long long)num_rois * pooled_h * pooled_w * pooled_h * pooled_w >= 0xFFFFFFFFLL)  // exact magic division by PH*PW
    return JTSM_ERR_UNSUPPORTED;
  if (workspace_bytes < L.total) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  if (num_rois > 0 && channels > 0 && (!grad_output || !rois)) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  const int cells = (int)cells_ll, hw = height * width, PP = pooled_h * pooled_w;
  AlignWorkspace ws = carve_align_workspace(workspace, batch, num_rois);
  char* p = (char*)(((uintptr_t)workspace + align_workspace_bytes(batch, num_rois) + 255) & ~(uintptr_t)255);
  unsigned* keys_a = (unsigned*)p; p += L.keys;
  unsigned* keys_b = (unsigned*)p; p += L.keys;
  unsigned long long* vals_a = (unsigned long long*)p; p += L.vals;
  unsigned long long* vals_b = (unsigned long long*)p; p += L.vals;
  int* first = (int*)p; p += L.first;
  float* counts = (float*)p; p += L.counts;
  float* gT = (float*)p; p += L.gt;
  void* cub_tmp = p;
  int rc = JTSM_OK;
  if (num_rois > 0) {
    rc = align_prep(rois, num_rois, batch, spatial_scale, pooled_h, pooled_w, sampling_ratio, aligned, ws, st);
    if (rc != JTSM_OK) return rc;
    const long long samples = (long long)num_rois * PP * sampling_ratio * sampling_ratio;
    align_det_taps_kernel<<<(unsigned)((samples + 255) / 256), 256, 0, st>>>(
        num_rois, height, width, pooled_h, pooled_w, sampling_ratio, ws, (unsigned)cells, keys_a, vals_a, counts);
    int end_bit = 1;
    while (end_bit < 32 && (1ull << end_bit) <= (unsigned long long)cells) ++end_bit;  // keys <= cells (sentinel)
    size_t tmp = L.cub;
    cub::DeviceRadixSort::SortPairs(cub_tmp, tmp, keys_a, keys_b, vals_a, vals_b, (long long)L.taps, 0, end_bit, st);
    dim3 tg((PP + 31) / 32, (channels + 31) / 32, num_rois);
    JTSM_DISPATCH_DTYPE(dtype, T, {
      align_det_transpose_kernel<T><<<tg, dim3(32, 32), 0, st>>>((const T*)grad_output, stride_n, stride_c, stride_h,
                                                                 stride_w, channels, pooled_h, pooled_w, gT);
    });
    count_launch(3);
  }
  align_det_bounds_kernel<<<(cells + 1 + 255) / 256, 256, 0, st>>>(keys_b, (long long)L.taps, cells, first);
  dim3 rg((cells + 31) / 32, (channels + 31) / 32);
  JTSM_DISPATCH_DTYPE(dtype, T, {
    align_det_reduce_kernel<T><<<rg, 1024, 0, st>>>(vals_b, first, counts, gT, channels, hw, cells, PP,
                                                     (T*)grad_input);
  });
  count_launch(2);
  return cuda_status();
}
