// RoIPool forward / backward for sm_100a.
//
// Reference arithmetic: torchvision.ops.roi_pool as constructed at
// projects/WSL/wsl/modeling/poolers.py:183-186; the same algorithm is restated
// in-tree at projects/WSL/wsl/layers/csrc/ROILoopPool/ROILoopPool_cpu.cpp:26-79
// (forward) and :83-123 (backward).
//
// B200 design (see DESIGN.md "RoIPool"):
//   * the feature map is NCHW and one channel plane (H*W*4 B = 30 KB at 75x100)
//     fits in shared memory, while every plane is revisited by all RoIs of its
//     image (~220 RoIs cover each cell at config 2).  So a CTA owns one
//     (image, channel) plane: it is read from HBM/L2 exactly once, coalesced,
//     and all RoI bins are then scanned out of shared memory.
//   * integer bin edges are channel independent: computed once per call into
//     a packed table (roi_bins.cuh) instead of per output element.
//   * threads map to consecutive (roi, bin) outputs, so for a fixed channel a
//     warp writes contiguous 4*P*P-byte runs of the [R, C, P, P] output.
#include <type_traits>

#include "roi_bins.cuh"
#include "roi_pool_rmq.cuh"
#include "roi_pool_rmqk.cuh"
#include "roi_pool_rmqp.cuh"

namespace jtsm {

// ---------------------------------------------------------------------------
// prep: bin tables + per-image RoI index range
// ---------------------------------------------------------------------------
// One CTA per kPrepRois RoIs: the first kPrepRois threads compute the bin-edge tables of one RoI each,
// then all threads turn them into the per-bin descriptors of the range-max-table kernel.
constexpr int kPrepRois = 32, kPrepThreads = 256;
__global__ void __launch_bounds__(kPrepThreads)
roi_pool_prep_kernel(const float* __restrict__ rois, int num_rois, int batch, int height, int width, float scale,
                     int PH, int PW, const float* __restrict__ row_scale, int want_desc, RmqBands bands,
                     PoolWorkspace ws) {
  const int r_base = blockIdx.x * kPrepRois;
  const int r = r_base + threadIdx.x;
  if (threadIdx.x < kPrepRois && r < num_rois) {
    const float* q = rois + (size_t)r * 5;
    const int b = (int)q[0];
    const RoiRect rc = roi_rect(q[1], q[2], q[3], q[4], scale);
    const int roi_w = max(rc.ew - rc.sw + 1, 1);
    const int roi_h = max(rc.eh - rc.sh + 1, 1);
    for (int p = 0; p < PH; ++p) ws.hbins[(size_t)r * PH + p] = bin_range(p, rc.sh, roi_h, PH, height);
    for (int p = 0; p < PW; ++p) ws.wbins[(size_t)r * PW + p] = bin_range(p, rc.sw, roi_w, PW, width);
    const bool ok = (b >= 0 && b < batch);
    ws.batch_idx[r] = ok ? b : -1;
    if (r > 0 && (int)rois[(size_t)(r - 1) * 5] > b) *ws.ungrouped = 1;
    if (want_desc) ws.rscale[r] = row_scale ? row_scale[r] : 1.f;
    if (ok) {
      atomicMin(ws.first + b, r);
      atomicMin(ws.neg_last + b, -r);
      if (want_desc && bands.nb > 1 && bands.nb <= kBandListBands) {
        // bins per (image, band): a bin belongs to the band of its first row (non-decreasing in the bin row), so a RoI's
        // bins of one band are a run of whole bin rows
        int run_band = -1, run_rows = 0;
        for (int p = 0; p <= PH; ++p) {
          const int bb = p < PH ? band_of(range_lo(ws.hbins[(size_t)r * PH + p]), bands) : -2;
          if (bb != run_band) {
            if (run_rows > 0) atomicAdd(ws.band_cnt + b * kBandListBands + run_band, run_rows * PW);
            run_band = bb, run_rows = 0;
          }
          ++run_rows;
        }
      }
    }
  }
  if (!want_desc) return;  // uniform
  __syncthreads();         // the CTA's bin tables are visible to all its threads
  // descriptors are relative to the bin's row band (one band = the whole plane when it fits shared memory)
  const int hw_t = bands.Hb * width, PP = PH * PW;
  const int n_local = min(kPrepRois, num_rois - r_base);
  for (int j = threadIdx.x; j < n_local * PP; j += kPrepThreads) {
    const int rl = j / PP, bin = j - rl * PP, ph = bin / PW, pw = bin - ph * PW;
    const int rr = r_base + rl;
    const uint32_t hb = ws.hbins[(size_t)rr * PH + ph];
    const int b = band_of(range_lo(hb), bands), row0 = b * bands.S;
    const bool fits = range_hi(hb) <= row0 + band_rows(b, height, bands) || range_hi(hb) <= range_lo(hb);
    const uint32_t hb_rel = hb - ((uint32_t)row0 | ((uint32_t)row0 << 16));
    const uint32_t d = fits ? rmq_bin_desc(hb_rel, ws.wbins[(size_t)rr * PW + pw], width, hw_t) : 0u;
    ws.bdesc[(size_t)rr * PP + bin] = d;
    ws.bband[(size_t)rr * PP + bin] = (unsigned char)b;
    if (d == 0u)  // ~2% of the bins; bit 30: taller than a band (RoI reaching far outside the map) -> global-memory scan
      ws.slow_list[atomicAdd(ws.slow_count, 1)] = (rr * PP + bin) | (fits ? 0 : 0x40000000);
  }
}

// ---------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------
// grid.x = batch * channels * splits ; block = kThreads
// dynamic smem: height * pitch floats
template <typename T, typename A, int kThreads>
__global__ void __launch_bounds__(kThreads)
roi_pool_fwd_kernel(const T* __restrict__ input, int channels, int height, int width, int pitch,
                    int PH, int PW, int splits, PoolWorkspace ws,
                    const float* __restrict__ row_scale, T* __restrict__ output,
                    A* __restrict__ argmax) {
  extern __shared__ float plane[];
  const int item = blockIdx.x;
  const int s = item % splits;
  const int nc = item / splits;
  const int n = nc / channels;
  const int c = nc - n * channels;

  const int first = ws.first[n];
  const int last = -ws.neg_last[n];
  if (first > last) return;  // image has no RoIs (uniform per CTA)
  const int len = last - first + 1;
  const int r0 = first + (int)(((long long)len * s) / splits);
  const int r1 = first + (int)(((long long)len * (s + 1)) / splits);
  if (r0 >= r1) return;

  // stage the plane: coalesced read, row pitch made odd for bank spread
  const T* src = input + (size_t)nc * height * width;
  for (int y = threadIdx.x / 32; y < height; y += kThreads / 32) {
    const T* row = src + (size_t)y * width;
    float* drow = plane + y * pitch;
    for (int x = threadIdx.x % 32; x < width; x += 32) drow[x] = Elem<T>::to_f(row[x]);
  }
  __syncthreads();

  const int PP = PH * PW;
  const int total = (r1 - r0) * PP;
  for (int i = threadIdx.x; i < total; i += kThreads) {
    const int rl = i / PP;
    const int bin = i - rl * PP;
    const int ph = bin / PW;
    const int pw = bin - ph * PW;
    const int r = r0 + rl;
    const size_t o = ((size_t)r * channels + c) * PP + bin;
    if (ws.batch_idx[r] != n) continue;  // only when RoIs are not grouped by image
    const uint32_t hb = __ldg(ws.hbins + (size_t)r * PH + ph);
    const uint32_t wb = __ldg(ws.wbins + (size_t)r * PW + pw);
    const int hs = range_lo(hb), he = range_hi(hb);
    const int wstart = range_lo(wb), wend = range_hi(wb);
    const bool empty = (he <= hs) || (wend <= wstart);
    float m = empty ? 0.f : -FLT_MAX;
    int mi = -1;
    for (int h = hs; h < he; ++h) {
      const float* prow = plane + h * pitch;
      for (int w = wstart; w < wend; ++w) {
        const float v = prow[w];
        if (v > m) {
          m = v;
          mi = h * width + w;
        }
      }
    }
    if (row_scale) m = __fmul_rn(m, row_scale[r]);
    output[o] = Elem<T>::from_f(m);
    argmax[o] = Arg<A>::enc(mi);
  }
}

// Fallback for planes that do not fit in shared memory: same mapping, reads the
// plane through L1/L2.
template <typename T, typename A, int kThreads>
__global__ void __launch_bounds__(kThreads)
roi_pool_fwd_gmem_kernel(const T* __restrict__ input, int channels, int height, int width,
                         int PH, int PW, int splits, PoolWorkspace ws,
                         const float* __restrict__ row_scale, T* __restrict__ output,
                         A* __restrict__ argmax) {
  const int item = blockIdx.x;
  const int s = item % splits;
  const int nc = item / splits;
  const int n = nc / channels;
  const int c = nc - n * channels;
  const int first = ws.first[n];
  const int last = -ws.neg_last[n];
  if (first > last) return;
  const int len = last - first + 1;
  const int r0 = first + (int)(((long long)len * s) / splits);
  const int r1 = first + (int)(((long long)len * (s + 1)) / splits);
  const T* src = input + (size_t)nc * height * width;
  const int PP = PH * PW;
  const int total = (r1 - r0) * PP;
  for (int i = threadIdx.x; i < total; i += kThreads) {
    const int rl = i / PP;
    const int bin = i - rl * PP;
    const int ph = bin / PW;
    const int pw = bin - ph * PW;
    const int r = r0 + rl;
    const size_t o = ((size_t)r * channels + c) * PP + bin;
    if (ws.batch_idx[r] != n) continue;
    const uint32_t hb = __ldg(ws.hbins + (size_t)r * PH + ph);
    const uint32_t wb = __ldg(ws.wbins + (size_t)r * PW + pw);
    const int hs = range_lo(hb), he = range_hi(hb);
    const int wstart = range_lo(wb), wend = range_hi(wb);
    const bool empty = (he <= hs) || (wend <= wstart);
    float m = empty ? 0.f : -FLT_MAX;
    int mi = -1;
    for (int h = hs; h < he; ++h) {
      const T* prow = src + (size_t)h * width;
      for (int w = wstart; w < wend; ++w) {
        const float v = Elem<T>::to_f(__ldg(prow + w));
        if (v > m) {
          m = v;
          mi = h * width + w;
        }
      }
    }
    if (row_scale) m = __fmul_rn(m, row_scale[r]);
    output[o] = Elem<T>::from_f(m);
    argmax[o] = Arg<A>::enc(mi);
  }
}

// ---------------------------------------------------------------------------
// backward, variant 0: one RED per output element straight into L2
// (grad_input fp32; 16-bit gradients accumulate in an fp32 workspace).
// ---------------------------------------------------------------------------
template <typename T, typename A>
__global__ void __launch_bounds__(256)
roi_pool_bwd_red_kernel(const T* __restrict__ grad_output, long long sn, long long sc, long long sh,
                        long long sw, const float* __restrict__ rois, const A* __restrict__ argmax,
                        int batch, int channels, int hw, int num_rois, int PH, int PW,
                        const float* __restrict__ row_scale, float* __restrict__ grad_acc) {
  const int PP = PH * PW;
  const long long total = (long long)num_rois * channels * PP;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int a = Arg<A>::dec(argmax[e]);
    if (a < 0) continue;
    const int bin = (int)(e % PP);
    const long long rc = e / PP;
    const int c = (int)(rc % channels);
    const int r = (int)(rc / channels);
    const int b = (int)__ldg(rois + (size_t)r * 5);
    if (b < 0 || b >= batch) continue;
    const int ph = bin / PW, pw = bin - ph * PW;
    float g = Elem<T>::to_f(grad_output[r * sn + c * sc + ph * sh + pw * sw]);
    if (row_scale) g = __fmul_rn(g, row_scale[r]);
    atomicAdd(grad_acc + ((size_t)b * channels + c) * hw + a, g);
  }
}


// ---------------------------------------------------------------------------
// backward, contiguous fast path: one CTA row per RoI, 4 consecutive outputs per thread
// (128-bit gradient load + 64/128-bit argmax load), no integer division (magic multiply),
// one RED.ADD.F32 per non-empty bin into the L2-resident gradient map.
// ---------------------------------------------------------------------------
template <typename A> struct Arg4;
template <> struct Arg4<unsigned short> {
  static __device__ __forceinline__ void load(const unsigned short* p, int (&a)[4]) {
    const uint2 v = *reinterpret_cast<const uint2*>(p);
    a[0] = Arg<unsigned short>::dec((unsigned short)(v.x & 0xFFFFu));
    a[1] = Arg<unsigned short>::dec((unsigned short)(v.x >> 16));
    a[2] = Arg<unsigned short>::dec((unsigned short)(v.y & 0xFFFFu));
    a[3] = Arg<unsigned short>::dec((unsigned short)(v.y >> 16));
  }
};
template <> struct Arg4<int> {
  static __device__ __forceinline__ void load(const int* p, int (&a)[4]) {
    const int4 v = *reinterpret_cast<const int4*>(p);
    a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w;
  }
};
template <typename T> struct Grad4;
template <> struct Grad4<float> {
  static __device__ __forceinline__ void load(const float* p, float (&g)[4]) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    g[0] = v.x; g[1] = v.y; g[2] = v.z; g[3] = v.w;
  }
};
template <> struct Grad4<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&g)[4]) {
    const uint2 v = *reinterpret_cast<const uint2*>(p);
    g[0] = __uint_as_float(v.x << 16); g[1] = __uint_as_float(v.x & 0xFFFF0000u);
    g[2] = __uint_as_float(v.y << 16); g[3] = __uint_as_float(v.y & 0xFFFF0000u);
  }
};
template <> struct Grad4<__half> {
  static __device__ __forceinline__ void load(const __half* p, float (&g)[4]) {
    const uint2 v = *reinterpret_cast<const uint2*>(p);
    const __half2 lo = *reinterpret_cast<const __half2*>(&v.x), hi = *reinterpret_cast<const __half2*>(&v.y);
    g[0] = __low2float(lo); g[1] = __high2float(lo); g[2] = __low2float(hi); g[3] = __high2float(hi);
  }
};

template <typename T, typename A>
__global__ void __launch_bounds__(256)
roi_pool_bwd_vec4_kernel(const T* __restrict__ grad_output, const float* __restrict__ rois,
                         const A* __restrict__ argmax, int batch, int channels, int hw, int PP,
                         unsigned magic_pp, const float* __restrict__ row_scale,
                         float* __restrict__ grad_acc) {
  const int r = blockIdx.x;
  const int b = (int)__ldg(rois + (size_t)r * 5);
  if (b < 0 || b >= batch) return;
  const float rs = row_scale ? __ldg(row_scale + r) : 1.f;
  const int cpp = channels * PP;
  const T* g_r = grad_output + (size_t)r * cpp;
  const A* a_r = argmax + (size_t)r * cpp;
  float* acc_b = grad_acc + (size_t)b * channels * hw;
  for (int i = (blockIdx.y * 256 + threadIdx.x) * 4; i < cpp; i += gridDim.y * 1024) {
    int a[4];
    float g[4];
    Arg4<A>::load(a_r + i, a);
    Grad4<T>::load(g_r + i, g);
    int cell[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = (int)__umulhi((unsigned)(i + j), magic_pp);
      cell[j] = a[j] >= 0 ? c * hw + a[j] : -1;
      g[j] = __fmul_rn(g[j], rs);
    }
    // neighbouring bins often share their arg-max cell (18% of adjacent pairs at config 2): fold
    // those in registers so the L2 sees one RED instead of two
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      if (cell[j] >= 0 && cell[j] == cell[j + 1]) {
        g[j + 1] = __fadd_rn(g[j + 1], g[j]);
        cell[j] = -1;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (cell[j] >= 0) atomicAdd(acc_b + cell[j], g[j]);
  }
}

template <typename T>
__global__ void convert_from_f32_kernel(const float* __restrict__ src, T* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = Elem<T>::from_f(src[i]);
}

}  // namespace jtsm

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
using namespace jtsm;

namespace {

constexpr int kFwdThreads = 256;
constexpr size_t kMaxPlaneSmem = 200 * 1024;

// The range-max-table kernel applies when plane + tables fit in shared memory (hw <= 9685, which also
// keeps every descriptor field inside its bits), the batch index fits 11 bits and output offsets 32 bits.
// Row bands for a plane (RmqBands, roi_bins.cuh): one band when plane + tables fit shared memory, else bands of the
// largest height whose plane + tables + staging buffer fit, overlapping by the tallest bin of a RoI inside the map.
// nb = 0: not feasible (very wide maps, where a band would be shorter than ~1.5 bins) -> direct-scan kernels.
// packed: the 16-bit kernel of roi_pool_rmqp.cuh (40 B per cell, no fp32 plane) instead of 24 B per cell.
RmqBands pick_bands(int height, int width, int PH, bool packed = false) {
  const long long hw = (long long)height * width;
  if (hw <= 0 || width > 4096) return {0, 0, 0};
  if (packed) {
    if (hw < (1 << 20) && rp_smem_core((int)hw) <= 227 * 1024) return {1, height, height};
    // 40 B of tables + 2 B of bulk-copy staging per cell: with only two or three bands the band's plane is worth
    // fetching behind the previous band's queries (the index kernel's many-band geometries prefer taller bands)
    int Hb = (int)((227 * 1024 - kRpHeader - 128) / (42 * (long long)width));
    while (Hb > 0 && rp_smem_core(Hb * width) + rk_stage_bytes(Hb * width, 2) > 227 * 1024) --Hb;
    const int O = (height + PH - 1) / PH + 1;
    const int S = Hb - O;
    if (Hb < 3 || S < 1 || 3 * S < Hb || (long long)Hb * width > 65535) return {0, 0, 0};
    const int nb = (height - O + S - 1) / S;
    if (nb < 1 || nb > 255) return {0, 0, 0};
    return {nb, S, Hb};
  }
  if (hw < (1 << 20) && rk_smem_core((int)hw) <= 227 * 1024) return {1, height, height};
  // bands as tall as plane + tables allow (24 B per cell): no staging buffer in banded mode -- fewer bands (fewer table
  // builds and descriptor sweeps per plane) are worth more than the bulk-copy fetch (150 x 200: 6 bands instead of 8)
  int Hb = (int)((227 * 1024 - kRkHeader - 64) / (24 * (long long)width));
  while (Hb > 0 && rk_smem_core(Hb * width) > 227 * 1024) --Hb;
  const int O = (height + PH - 1) / PH + 1;
  const int S = Hb - O;
  if (Hb < 3 || S < 1 || 3 * S < Hb) return {0, 0, 0};
  const int nb = (height - O + S - 1) / S;
  if (nb < 1 || nb > 255) return {0, 0, 0};
  return {nb, S, Hb};
}

bool rmq_geometry_ok(int batch, int channels, int height, int width, int num_rois, int PH, int PW) {
  const long long hw = (long long)height * width;
  return pick_bands(height, width, PH).nb > 0 &&
         (long long)num_rois * channels * PH * PW < 0x7FFFFFFFLL &&
         (long long)num_rois * PH * PW * PH * PW < 0xFFFFFFFFLL &&  // exact magic division of a bin id by PH*PW
         (long long)num_rois * 64 < 0x7FFFFFFFLL &&                // chunk bounds in 32 bits (chunks <= 64)
         (long long)batch * channels * hw * 4 < 0xFFFFFFF0LL;      // plane byte offsets in 32 bits
}
bool rmq_eligible(int batch, int channels, int height, int width, int num_rois, int PH, int PW) {
  const int algo = tuning("roi_pool_fwd_algo", 0);  // 0 auto, 1 direct smem scan, 2 gmem scan, 3 rmq
  return (algo == 0 || algo == 3) && !tuning("roi_pool_force_gmem", 0) &&
         rmq_geometry_ok(batch, channels, height, width, num_rois, PH, PW);
}

int pick_splits(int planes, int num_rois, int ctas_per_sm) {
  // enough CTAs for ~4 waves, but keep >= 64 RoIs per CTA so the plane load amortises
  const int forced = tuning("roi_pool_splits", 0);
  if (forced > 0) return forced;
  const long long target = 4LL * num_sms() * ctas_per_sm;
  long long s = (target + planes - 1) / planes;
  const long long cap = num_rois / 64 > 1 ? num_rois / 64 : 1;
  if (s > cap) s = cap;
  if (s < 1) s = 1;
  return (int)s;
}

// range-max-table kernel: plane + 10 index tables in shared memory, persistent CTAs.  kMoi: MOIPool mode (MoiAux).
template <typename T, typename A, bool kMoi>
int launch_rmq(const void* input, int batch, int channels, int height, int width, int num_rois, int PH, int PW,
               const PoolWorkspace& ws, const MoiAux& moi, void* output, void* argmax, cudaStream_t st) {
  const RmqBands bands = pick_bands(height, width, PH);
  if (bands.nb < 1 || (kMoi && bands.nb != 1)) return JTSM_ERR_UNSUPPORTED;
  const int real_planes = batch * channels;
  const int planes = real_planes * bands.nb;  // virtual planes: (band, plane)
  const int hw = bands.Hb * width;            // cells of a band
  {
    // table format: index tables (8 look-ups per bin, LSU bound) by default; keyed entries (5 look-ups, ALU bound and
    // slower on B200: profiles/r2_roi_pool_fwd_keyed.md) on request.  Planes arrive by bulk copy (TMA 1-D) into a raw
    // staging buffer when one fits next to plane + tables.
    const bool keyed = tuning("roi_pool_fwd_keys", 0) != 0;
    const size_t stage = rk_stage_bytes(hw, (int)sizeof(T));
    const int staged = rk_smem_core(hw) + stage <= 227 * 1024 && tuning("roi_pool_fwd_tma", 1);
    const size_t smem = rk_smem_core(hw) + (staged ? stage : 0);
    // kernel instantiation: table format x banded (the masked pools are single band only)
    auto kern = roi_pool_fwd_rmq_kernel<T, A, kMoi, false, false>;
    if (keyed && bands.nb > 1) kern = roi_pool_fwd_rmq_kernel<T, A, kMoi, true, !kMoi>;
    else if (keyed) kern = roi_pool_fwd_rmq_kernel<T, A, kMoi, true, false>;
    else if (bands.nb > 1) kern = roi_pool_fwd_rmq_kernel<T, A, kMoi, false, !kMoi>;
    // per call: the attribute is per device, and one process may drive several GPUs
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    const int per_sm = (int)((227 * 1024) / (smem + 1024)) >= 2 ? 2 : 1;  // 1024-thread CTAs
    const int max_grid = num_sms() * per_sm;
    // RoI chunks per plane: the planes of the last, partial round (planes % grid) are cut into chunks and dealt to
    // all CTAs, so the round costs ceil(tail_planes * chunks / grid) / chunks plane-times; every extra chunk adds
    // a little per-item overhead (and a table build when a CTA's range crosses planes).  Pick the cheapest.
    int chunks = tuning("roi_pool_splits", 0);
    if (chunks <= 0) {
      const int cap = num_rois / 64 > 1 ? (num_rois / 64 < 64 ? num_rois / 64 : 64) : 1;
      const int tail = planes % max_grid;
      double best = 1e30;
      chunks = 1;
      for (int c = 1; c <= cap && tail > 0; ++c) {
        const double cost = (double)((tail * c + max_grid - 1) / max_grid) / c + 0.01 * c;
        if (cost < best - 1e-9) best = cost, chunks = c;
      }
    }
    if (chunks > 64) chunks = 64;
    const int items = planes * chunks;
    const int grid = items < max_grid ? items : max_grid;
    const unsigned long long in_bytes = (unsigned long long)real_planes * height * width * sizeof(T);
    // bit 1: the row-mapped main loop (11 % fewer instructions, same time: the kernel is LSU bound at c2 -- opt-in,
    // profiles/r2o_roi_pool_pack16.md)
    // bit 2: banded planes walk the per-band RoI lists instead of filtering every bin by its band
    const int flags = staged | (tuning("roi_pool_fwd_rows", 0) ? 2 : 0) |
                      (bands.nb > 1 && bands.nb <= kBandListBands && tuning("roi_pool_fwd_bandlist", 1) ? 4 : 0);
    kern<<<grid, kRmqThreads, smem, st>>>((const T*)input, in_bytes, real_planes, channels, height, width, PH, PW, chunks,
                                          items, flags, bands, ws, moi, (T*)output, (A*)argmax);
    count_launch();
    return cuda_status();
  }
}

// 16-bit feature maps: packed (value, index) tables (roi_pool_rmqp.cuh) when the band geometry allows it.
bool rmqp_eligible(int dtype, int batch, int channels, int height, int width, int num_rois, int PH, int PW) {
  return dtype != JTSM_F32 && tuning("roi_pool_fwd_pack16", 1) && !tuning("roi_pool_fwd_keys", 0) &&
         rmq_eligible(batch, channels, height, width, num_rois, PH, PW) &&
         // measured (profiles/r2o_roi_pool_pack16.md): wins 15 % while the plane fits ONE band (<= 5.8 K cells: stride-16 /
         // stride-32 maps); with row bands the band filter of the main loop costs more than the saved look-ups
         // (75 x 100: 318 us against 228 us), so banded geometries keep the index tables unless forced (knob = 2)
         (pick_bands(height, width, PH, true).nb == 1 ||
          (tuning("roi_pool_fwd_pack16", 1) == 2 && pick_bands(height, width, PH, true).nb > 1));
}

template <typename T, typename A>
int launch_rmqp(const void* input, int batch, int channels, int height, int width, int num_rois, int PH, int PW,
                const PoolWorkspace& ws, void* output, void* argmax, cudaStream_t st) {
  const RmqBands bands = pick_bands(height, width, PH, true);
  if (bands.nb < 1) return JTSM_ERR_UNSUPPORTED;
  const int real_planes = batch * channels;
  const int planes = real_planes * bands.nb;
  const int hw = bands.Hb * width;
  const size_t stage = rk_stage_bytes(hw, (int)sizeof(T));
  const int staged = rp_smem_core(hw) + stage <= 227 * 1024 && tuning("roi_pool_fwd_tma", 1);
  const size_t smem = rp_smem_core(hw) + (staged ? stage : 0);
  auto kern = roi_pool_fwd_rmqp_kernel<T, A, false>;
  if (bands.nb > 1) kern = roi_pool_fwd_rmqp_kernel<T, A, true>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);  // per call: per device
  const int per_sm = (int)((227 * 1024) / (smem + 1024)) >= 2 ? 2 : 1;
  const int max_grid = num_sms() * per_sm;
  int chunks = tuning("roi_pool_splits", 0);
  if (chunks <= 0) {
    const int cap = num_rois / 64 > 1 ? (num_rois / 64 < 64 ? num_rois / 64 : 64) : 1;
    const int tail = planes % max_grid;
    double best = 1e30;
    chunks = 1;
    for (int c = 1; c <= cap && tail > 0; ++c) {
      const double cost = (double)((tail * c + max_grid - 1) / max_grid) / c + 0.01 * c;
      if (cost < best - 1e-9) best = cost, chunks = c;
    }
  }
  if (chunks > 64) chunks = 64;
  const int items = planes * chunks;
  const int grid = items < max_grid ? items : max_grid;
  const unsigned long long in_bytes = (unsigned long long)real_planes * height * width * sizeof(T);
  const int flags = staged | (bands.nb > 1 && bands.nb <= kBandListBands && tuning("roi_pool_fwd_bandlist", 1) ? 4 : 0);
  kern<<<grid, kRmqThreads, smem, st>>>((const T*)input, in_bytes, real_planes, channels, height, width, PH, PW, chunks,
                                        items, flags, bands, ws, (T*)output, (A*)argmax);
  count_launch();
  return cuda_status();
}
template <typename T, typename A>
int launch_roi_pool_fwd(const void* input, int batch, int channels, int height, int width,
                        int num_rois, int PH, int PW, const PoolWorkspace& ws,
                        const float* row_scale, void* output, void* argmax, cudaStream_t st) {
  const int planes = batch * channels;
  const int algo = tuning("roi_pool_fwd_algo", 0);
  if constexpr (sizeof(T) == 2) {
    const int dt = std::is_same<T, __half>::value ? JTSM_F16 : JTSM_BF16;
    if (rmqp_eligible(dt, batch, channels, height, width, num_rois, PH, PW))
      return launch_rmqp<T, A>(input, batch, channels, height, width, num_rois, PH, PW, ws, output, argmax, st);
  }
  if (rmq_eligible(batch, channels, height, width, num_rois, PH, PW))
    return launch_rmq<T, A, false>(input, batch, channels, height, width, num_rois, PH, PW, ws, MoiAux{}, output,
                                   argmax, st);
  const int pitch = (width & 1) ? width : width + 1;
  const size_t smem = (size_t)height * pitch * sizeof(float);
  if (smem <= kMaxPlaneSmem && algo != 2 && !tuning("roi_pool_force_gmem", 0)) {
    auto k = roi_pool_fwd_kernel<T, A, kFwdThreads>;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxPlaneSmem);  // per device
    int per_sm = (int)((227 * 1024) / (smem + 1024));
    per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
    const int splits = pick_splits(planes, num_rois, per_sm);
    k<<<planes * splits, kFwdThreads, smem, st>>>((const T*)input, channels, height, width, pitch, PH,
                                                   PW, splits, ws, row_scale, (T*)output, (A*)argmax);
  } else {
    const int splits = pick_splits(planes, num_rois, 8);
    roi_pool_fwd_gmem_kernel<T, A, kFwdThreads><<<planes * splits, kFwdThreads, 0, st>>>(
        (const T*)input, channels, height, width, PH, PW, splits, ws, row_scale, (T*)output,
        (A*)argmax);
  }
  count_launch();
  return cuda_status();
}

// band lists, pass 2: offsets of the (image, band) segments, then every RoI writes the ids of its bins into its bands'
// segments (a run of whole bin rows per band; the order of the runs inside a segment is whatever the atomics give --
// every bin is written exactly once by the pooling kernel either way)
__global__ void band_scan_kernel(int slots, PoolWorkspace ws) {
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int i = 0; i < slots; ++i) {
      ws.band_off[i] = acc;
      acc += ws.band_cnt[i];
      ws.band_cur[i] = 0;
    }
    ws.band_off[slots] = acc;
  }
}
__global__ void band_fill_kernel(int num_rois, int PH, int PW, RmqBands bands, PoolWorkspace ws) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= num_rois) return;
  const int b = ws.batch_idx[r];
  if (b < 0) return;
  int run_band = -1, run_rows = 0, run_p0 = 0;
  for (int p = 0; p <= PH; ++p) {
    const int bb = p < PH ? band_of(range_lo(ws.hbins[(size_t)r * PH + p]), bands) : -2;
    if (bb != run_band) {
      if (run_rows > 0) {
        const int slot = b * kBandListBands + run_band, n = run_rows * PW;
        int* dst = ws.band_list + ws.band_off[slot] + atomicAdd(ws.band_cur + slot, n);
        const int id0 = r * PH * PW + run_p0 * PW;
        for (int k = 0; k < n; ++k) dst[k] = id0 + k;
      }
      run_band = bb, run_rows = 0, run_p0 = p;
    }
    ++run_rows;
  }
}

int run_prep(const float* rois, int num_rois, int batch, int channels, int height, int width, float scale,
             int PH, int PW, const float* row_scale, const PoolWorkspace& ws, cudaStream_t st, int force_desc = 0,
             bool packed = false) {
  cudaMemsetAsync(ws.first, 0x7f, (size_t)batch * 2 * sizeof(int), st);
  cudaMemsetAsync(ws.ungrouped, 0, 2 * sizeof(int), st);  // ungrouped flag + slow-bin counter
  cudaMemsetAsync(ws.band_cnt, 0, (size_t)batch * kBandListBands * sizeof(int), st);
  const int want_desc = force_desc || rmq_eligible(batch, channels, height, width, num_rois, PH, PW);
  RmqBands bands = pick_bands(height, width, PH, packed);  // the geometry of the kernel that will read the descriptors
  if (bands.nb < 1) bands = {1, height, height};  // descriptors unused then (direct-scan kernels)
  roi_pool_prep_kernel<<<(num_rois + kPrepRois - 1) / kPrepRois, kPrepThreads, 0, st>>>(
      rois, num_rois, batch, height, width, scale, PH, PW, row_scale, want_desc, bands, ws);
  count_launch();
  if (want_desc && bands.nb > 1 && bands.nb <= kBandListBands && num_rois > 0) {
    band_scan_kernel<<<1, 32, 0, st>>>(batch * kBandListBands, ws);
    band_fill_kernel<<<(num_rois + 255) / 256, 256, 0, st>>>(num_rois, PH, PW, bands, ws);
    count_launch(2);
  }
  return cuda_status();
}

}  // namespace

// ---- entry points for MOIPool's range-max mode (wsl_pool.cu) ----
namespace jtsm {
bool moi_rmq_eligible(int batch, int channels, int height, int width, int num_rois, int PH, int PW) {
  // the masked pools keep whole planes in shared memory (their visit-list kernels are not banded)
  return pick_bands(height, width, PH).nb == 1 && rmq_geometry_ok(batch, channels, height, width, num_rois, PH, PW);
}
int moi_rmq_prep(const float* rois, int num_rois, int batch, int channels, int height, int width, float scale, int PH,
                 int PW, const float* row_scale, const PoolWorkspace& ws, cudaStream_t st) {
  return run_prep(rois, num_rois, batch, channels, height, width, scale, PH, PW, row_scale, ws, st, 1);
}
int moi_rmq_launch(int dtype, const void* input, int batch, int channels, int height, int width, int num_rois, int PH,
                   int PW, const PoolWorkspace& ws, const MoiAux& moi, void* output, int* argmax, cudaStream_t st) {
  int rc = JTSM_OK;
  JTSM_DISPATCH_DTYPE(dtype, T, {
    rc = launch_rmq<T, int, true>(input, batch, channels, height, width, num_rois, PH, PW, ws, moi, output, argmax, st);
  });
  return rc;
}
}  // namespace jtsm

extern "C" size_t jtsm_roi_pool_workspace_bytes(int batch, int num_rois, int pooled_h, int pooled_w) {
  if (batch < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0) return 0;
  return pool_workspace_bytes(batch, num_rois, pooled_h, pooled_w, 1);
}

extern "C" int jtsm_roi_pool_forward(const void* input, const float* rois, int dtype, int batch,
                                     int channels, int height, int width, int num_rois,
                                     float spatial_scale, int pooled_h, int pooled_w,
                                     const float* row_scale, void* output, void* argmax,
                                     int argmax_dtype, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 ||
      pooled_w <= 0 || pooled_h > 64 || pooled_w > 64)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (height > 32767 || width > 32767) return JTSM_ERR_UNSUPPORTED;
  if (argmax_dtype == JTSM_ARGMAX_U16 && (long long)height * width >= 65535) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois == 0 || channels == 0) return JTSM_OK;  // nothing to write
  if (!input || !rois || !output || !argmax || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (batch == 0 || height == 0 || width == 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_roi_pool_workspace_bytes(batch, num_rois, pooled_h, pooled_w))
    return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  PoolWorkspace ws = carve_pool_workspace(workspace, batch, num_rois, pooled_h, pooled_w, 1);
  const bool packed = rmqp_eligible(dtype, batch, channels, height, width, num_rois, pooled_h, pooled_w);
  int rc = run_prep(rois, num_rois, batch, channels, height, width, spatial_scale, pooled_h, pooled_w, row_scale,
                    ws, st, 0, packed);
  if (rc != JTSM_OK) return rc;
  JTSM_DISPATCH_DTYPE(dtype, T, JTSM_DISPATCH_ARGMAX(argmax_dtype, A, {
                        rc = launch_roi_pool_fwd<T, A>(input, batch, channels, height, width,
                                                       num_rois, pooled_h, pooled_w, ws, row_scale,
                                                       output, argmax, st);
                      }));
  return rc;
}

// 7 x 7 bins, contiguous gradient, channels % 32 == 0: a lane owns ALL 49 bins of one (RoI, channel), so every pair of
// neighbouring bins that picked the same cell (right, below, and the two diagonals below: 33 % of the elements at config
// 2, against the 14 % the 4-wide kernel can see) folds in registers before the reduction; the warp's 32 channels are one
// contiguous 32 x 49 run of the gradient, staged through shared memory with coalesced loads (lane stride 49 words:
// conflict free).  The bound of the backward is the number of REDs, nothing else.
constexpr int kFoldWarps = 8;
template <typename T, typename A>
__global__ void __launch_bounds__(kFoldWarps * 32, 2)
roi_pool_bwd_fold7_kernel(const T* __restrict__ grad_output, const float* __restrict__ rois,
                          const A* __restrict__ argmax, int batch, int channels, int hw, int num_rois, int super,
                          const float* __restrict__ row_scale, float* __restrict__ grad_acc) {
  extern __shared__ __align__(16) unsigned char fold_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* sg = (float*)fold_smem + warp * (32 * 49 * 2);
  int* sa = (int*)(sg + 32 * 49);
  const int groups = channels >> 5;
  const long long units = (long long)num_rois * groups;
  for (long long u = (long long)blockIdx.x * kFoldWarps + warp; u < units; u += (long long)gridDim.x * kFoldWarps) {
    // Order of the units: RoI-major inside a super-group of `super` channel groups whose accumulation planes fit L2
    // together.  Spreading the warps in flight over many planes keeps concurrent REDs off the same cells (config 1:
    // 181 us against 202 us channel-group-major); bounding the spread keeps the planes L2 resident when the whole
    // gradient map is not (config 3, 123 MB: 6.7 ms RoI-major over all 2048 channels against 3.2 ms).
    const long long per_super = (long long)num_rois * super;
    const int sgi = (int)(u / per_super);
    const long long rem = u - sgi * per_super;
    const int gs = min(super, groups - sgi * super);
    const int r = (int)(rem / gs), cg = sgi * super + (int)(rem % gs);
    const int b = (int)__ldg(rois + (size_t)r * 5);
    if (b < 0 || b >= batch) continue;  // warp uniform
    const float rs = row_scale ? __ldg(row_scale + r) : 1.f;
    const size_t base = ((size_t)r * channels + (size_t)cg * 32) * 49;
    const T* g_r = grad_output + base;
    const A* a_r = argmax + base;
#pragma unroll 7
    for (int k = 0; k < 49; ++k) {
      sg[k * 32 + lane] = Elem<T>::to_f(g_r[k * 32 + lane]);
      sa[k * 32 + lane] = Arg<A>::dec(a_r[k * 32 + lane]);
    }
    __syncwarp();
    float g[49];
    int a[49];
#pragma unroll
    for (int k = 0; k < 49; ++k) {
      g[k] = __fmul_rn(sg[lane * 49 + k], rs);
      a[k] = sa[lane * 49 + k];
    }
    __syncwarp();
    float* acc = grad_acc + ((size_t)b * channels + (size_t)cg * 32 + lane) * hw;
#pragma unroll
    for (int k = 0; k < 49; ++k) {
      const int ph = k / 7, pw = k % 7;
      bool done = a[k] < 0;
      if (pw < 6) {
        const bool m = !done && a[k] == a[k + 1];
        g[k + 1] = __fadd_rn(g[k + 1], m ? g[k] : 0.f);
        done |= m;
      }
      if (ph < 6) {
        if (pw > 0) {
          const bool m = !done && a[k] == a[k + 6];
          g[k + 6] = __fadd_rn(g[k + 6], m ? g[k] : 0.f);
          done |= m;
        }
        {
          const bool m = !done && a[k] == a[k + 7];
          g[k + 7] = __fadd_rn(g[k + 7], m ? g[k] : 0.f);
          done |= m;
        }
        if (pw < 6) {
          const bool m = !done && a[k] == a[k + 8];
          g[k + 8] = __fadd_rn(g[k + 8], m ? g[k] : 0.f);
          done |= m;
        }
      }
      if (!done) {
        atomicAdd(acc + a[k], g[k]);
      }
    }
  }
}

extern "C" size_t jtsm_roi_pool_backward_workspace_bytes(int dtype, int batch, int channels,
                                                         int height, int width) {
  if (dtype == JTSM_F32) return 256;
  return (size_t)batch * channels * height * width * sizeof(float) + 256;
}

extern "C" int jtsm_roi_pool_backward(const void* grad_output, long long stride_n,
                                      long long stride_c, long long stride_h, long long stride_w,
                                      const float* rois, const void* argmax, int argmax_dtype,
                                      int dtype, int batch, int channels, int height, int width,
                                      int num_rois, int pooled_h, int pooled_w,
                                      const float* row_scale, void* grad_input, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 ||
      pooled_w <= 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  const size_t n_in = (size_t)batch * channels * height * width;
  if (n_in == 0) return JTSM_OK;
  if (!grad_input) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esz = dtype == JTSM_F32 ? 4 : 2;
  float* acc = (float*)grad_input;
  if (dtype != JTSM_F32) {
    if (!workspace || workspace_bytes < jtsm_roi_pool_backward_workspace_bytes(dtype, batch, channels, height, width))
      return JTSM_ERR_WORKSPACE_TOO_SMALL;
    acc = (float*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  }
  cudaMemsetAsync(acc, 0, n_in * sizeof(float), st);
  (void)esz;
  const long long total = (long long)num_rois * channels * pooled_h * pooled_w;
  int rc = JTSM_OK;
  if (total > 0) {
    if (!grad_output || !rois || !argmax) return JTSM_ERR_INVALID_ARGUMENT;
    const int PP = pooled_h * pooled_w;
    const long long cpp = (long long)channels * PP;
    const bool contiguous = stride_w == 1 && stride_h == pooled_w && stride_c == PP && stride_n == cpp;
    const bool vec_ok = contiguous && cpp % 4 == 0 && cpp < (1ll << 31) && cpp * PP < (1ll << 32) &&  // exact magic division
                        
                        (long long)channels * height * width < (1ll << 31) &&
                        ((uintptr_t)grad_output % 16 == 0) && ((uintptr_t)argmax % 16 == 0) &&
                        tuning("roi_pool_bwd_algo", 0) != 1;
    const int bwd_algo = tuning("roi_pool_bwd_algo", 0);
    const bool fold_ok = contiguous && pooled_h == 7 && pooled_w == 7 && channels % 32 == 0 && bwd_algo != 1 &&
                         bwd_algo != 3 && (long long)channels * height * width < (1ll << 31);
    if (fold_ok) {
      const size_t smem = (size_t)kFoldWarps * 32 * 49 * 2 * sizeof(float);
      const long long units = (long long)num_rois * (channels / 32);
      long long blocks = (units + kFoldWarps - 1) / kFoldWarps;
      const long long cap = (long long)num_sms() * 2 * 8;
      if (blocks > cap) blocks = cap;
      // how many channel groups the warps in flight spread over (measured, profiles/r2j_roi_pool_bwd_fold.md): all of
      // them while the whole fp32 gradient map is small (<= 16 MB: configs 1, 2), else ~4 MB of planes (config 3)
      const long long group_bytes = (long long)batch * height * width * 4 * 32;
      const int knob_mb = tuning("roi_pool_bwd_super_mb", 0);
      long long budget = knob_mb > 0 ? ((long long)knob_mb << 20)
                                     : ((long long)n_in * 4 <= (16ll << 20) ? (long long)n_in * 4 : (4ll << 20));
      long long sup = budget / (group_bytes > 0 ? group_bytes : 1);
      const int super = (int)(sup < 1 ? 1 : (sup > channels / 32 ? channels / 32 : sup));
      JTSM_DISPATCH_DTYPE(dtype, T, JTSM_DISPATCH_ARGMAX(argmax_dtype, A, {
                            cudaFuncSetAttribute(roi_pool_bwd_fold7_kernel<T, A>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                            roi_pool_bwd_fold7_kernel<T, A><<<(int)blocks, kFoldWarps * 32, smem, st>>>(
                                (const T*)grad_output, rois, (const A*)argmax, batch, channels, height * width,
                                num_rois, super, row_scale, acc);
                          }));
    } else if (vec_ok) {
      const unsigned magic_pp = (unsigned)((0x100000000ull + PP - 1) / PP);
      int ysplit = (int)((cpp + 1023) / 1024);
      const int want = (4 * num_sms() * 8 + num_rois - 1) / num_rois;  // enough CTAs when R is small
      if (ysplit > want) ysplit = want < 1 ? 1 : want;
      dim3 grid(num_rois, ysplit);
      JTSM_DISPATCH_DTYPE(dtype, T, JTSM_DISPATCH_ARGMAX(argmax_dtype, A, {
                            roi_pool_bwd_vec4_kernel<T, A><<<grid, 256, 0, st>>>(
                                (const T*)grad_output, rois, (const A*)argmax, batch, channels,
                                height * width, PP, magic_pp, row_scale, acc);
                          }));
    } else {
      long long blocks = (total + 255) / 256;
      const long long cap = (long long)num_sms() * 8 * 4;
      if (blocks > cap) blocks = cap;
      JTSM_DISPATCH_DTYPE(dtype, T, JTSM_DISPATCH_ARGMAX(argmax_dtype, A, {
                            roi_pool_bwd_red_kernel<T, A><<<(int)blocks, 256, 0, st>>>(
                                (const T*)grad_output, stride_n, stride_c, stride_h, stride_w, rois,
                                (const A*)argmax, batch, channels, height * width, num_rois, pooled_h,
                                pooled_w, row_scale, acc);
                          }));
    }
    count_launch();
    rc = cuda_status();
    if (rc != JTSM_OK) return rc;
  }
  if (dtype == JTSM_BF16) {
    convert_from_f32_kernel<__nv_bfloat16><<<num_sms() * 8, 256, 0, st>>>(acc, (__nv_bfloat16*)grad_input, n_in);
    count_launch();
  } else if (dtype == JTSM_F16) {
    convert_from_f32_kernel<__half><<<num_sms() * 8, 256, 0, st>>>(acc, (__half*)grad_input, n_in);
    count_launch();
  }
  return cuda_status();
}
