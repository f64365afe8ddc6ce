// The WSL project's own pooling ops for sm_100a: ROILoopPool (RoI / frame / context triple max-pool)
// and MOIPool (superpixel-masked RoI max-pool).
//
// Reference (CUDA only; the CPU entry points raise, ROILoopPool.h:62, MOIPool.h:74):
//   projects/WSL/wsl/layers/csrc/ROILoopPool/ROILoopPool_cuda.cu:10-204 (forward), :207-248 (backward),
//     host wrappers :252-386 (context_ratio = 1.8 at :309);  Python: wsl/layers/roi_loop_pool.py:9-58
//   projects/WSL/wsl/layers/csrc/MOIPool/MOIPool_cuda.cu:139-215 (MoIForward mask), :218-294
//     (MoIPoolForward2), :297-338 (backward);  Python: wsl/layers/moi_pool.py:10-88
//
// Design: channel-independent geometry (integer rectangles, bin edges, the superpixel mask) is
// computed once per call; the max-pool itself runs one CTA per (image, channel) plane with the plane
// staged in shared memory (read from HBM once), threads over (RoI, bin).  Backward of both ops is the
// shared max-pool backward (jtsm_roi_pool_backward) over the saved argmax.
#include "roi_bins.cuh"
#include "roi_pool_rmq.cuh"

namespace jtsm {

// range-max mode for MOIPool (roi_pool.cu)
bool moi_rmq_eligible(int batch, int channels, int height, int width, int num_rois, int PH, int PW);
int moi_rmq_prep(const float* rois, int num_rois, int batch, int channels, int height, int width, float scale, int PH,
                 int PW, const float* row_scale, const PoolWorkspace& ws, cudaStream_t st);
int moi_rmq_launch(int dtype, const void* input, int batch, int channels, int height, int width, int num_rois, int PH,
                   int PW, const PoolWorkspace& ws, const MoiAux& moi, void* output, int* argmax, cudaStream_t st);

struct LoopRects {  // integer rectangles of one RoI (ROILoopPool_cuda.cu:35-104,146-170)
  int in0[4];       // inner rect (sw, sh, ew, eh) excluded from the frame pool
  int in1[4];       // RoI rect excluded from the context pool
};

__device__ __forceinline__ float clampf_(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// prep: bin tables for set 0 (RoI rect) and set 1 (outer rect) + exclusion rectangles
__global__ void loop_pool_prep_kernel(const float* __restrict__ rois, int num_rois, int batch, int H, int W,
                                      float scale, float ratio, int PH, int PW, PoolWorkspace ws,
                                      LoopRects* __restrict__ rects) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= num_rois) return;
  const float* q = rois + (size_t)r * 5;
  const int b = (int)q[0];
  const float x1 = q[1], y1 = q[2], x2 = q[3], y2 = q[4];
  const float rw = __fsub_rn(x2, x1), rh = __fsub_rn(y2, y1);
  const float iw = __fdiv_rn(rw, ratio), ih = __fdiv_rn(rh, ratio);
  const float ow = __fmul_rn(rw, ratio), oh = __fmul_rn(rh, ratio);
  const float irw = __fsub_rn(rw, iw), irh = __fsub_rn(rh, ih);
  const float orw = __fsub_rn(ow, rw), orh = __fsub_rn(oh, rh);
  const float lim_w = (float)(1.0 * W / (double)scale), lim_h = (float)(1.0 * H / (double)scale);
  const float x1i = clampf_(__fadd_rn(x1, __fdiv_rn(irw, 2.f)), 0.f, lim_w);
  const float y1i = clampf_(__fadd_rn(y1, __fdiv_rn(irh, 2.f)), 0.f, lim_h);
  const float x2i = clampf_(__fsub_rn(x2, __fdiv_rn(irw, 2.f)), 0.f, lim_w);
  const float y2i = clampf_(__fsub_rn(y2, __fdiv_rn(irh, 2.f)), 0.f, lim_h);
  const float x1o = clampf_(__fsub_rn(x1, __fdiv_rn(orw, 2.f)), 0.f, lim_w);
  const float y1o = clampf_(__fsub_rn(y1, __fdiv_rn(orh, 2.f)), 0.f, lim_h);
  const float x2o = clampf_(__fadd_rn(x2, __fdiv_rn(orw, 2.f)), 0.f, lim_w);
  const float y2o = clampf_(__fadd_rn(y2, __fdiv_rn(orh, 2.f)), 0.f, lim_h);
  const RoiRect roi = roi_rect(x1, y1, x2, y2, scale);
  const RoiRect inner = roi_rect(x1i, y1i, x2i, y2i, scale);
  const RoiRect outer = roi_rect(x1o, y1o, x2o, y2o, scale);
  LoopRects lr;
  lr.in0[0] = inner.sw; lr.in0[1] = inner.sh; lr.in0[2] = inner.ew; lr.in0[3] = inner.eh;
  lr.in1[0] = roi.sw; lr.in1[1] = roi.sh; lr.in1[2] = roi.ew; lr.in1[3] = roi.eh;
  rects[r] = lr;
  const size_t set_h = (size_t)num_rois * PH, set_w = (size_t)num_rois * PW;
  {
    const int w_ = max(roi.ew - roi.sw + 1, 1), h_ = max(roi.eh - roi.sh + 1, 1);
    for (int p = 0; p < PH; ++p) ws.hbins[(size_t)r * PH + p] = bin_range(p, roi.sh, h_, PH, H);
    for (int p = 0; p < PW; ++p) ws.wbins[(size_t)r * PW + p] = bin_range(p, roi.sw, w_, PW, W);
  }
  {
    const int w_ = max(outer.ew - outer.sw + 1, 1), h_ = max(outer.eh - outer.sh + 1, 1);
    for (int p = 0; p < PH; ++p) ws.hbins[set_h + (size_t)r * PH + p] = bin_range(p, outer.sh, h_, PH, H);
    for (int p = 0; p < PW; ++p) ws.wbins[set_w + (size_t)r * PW + p] = bin_range(p, outer.sw, w_, PW, W);
  }
  const bool ok = b >= 0 && b < batch;
  ws.batch_idx[r] = ok ? b : -1;
  if (ok) {
    atomicMin(ws.first + b, r);
    atomicMin(ws.neg_last + b, -r);
  }
}

constexpr int kWslThreads = 256;

// value fetch from the staged plane or straight from global memory
template <typename T, bool kSmem>
__device__ __forceinline__ float plane_at(const float* plane_s, const T* src, int idx) {
  return kSmem ? plane_s[idx] : Elem<T>::to_f(__ldg(src + idx));
}

template <typename T, bool kSmem>
__global__ void __launch_bounds__(kWslThreads)
loop_pool_fwd_kernel(const T* __restrict__ input, int channels, int H, int W, int PH, int PW, int splits,
                     int num_rois, PoolWorkspace ws, const LoopRects* __restrict__ rects,
                     T* __restrict__ output, int* __restrict__ argmax) {
  extern __shared__ float plane_s[];
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
  if (r0 >= r1) return;
  const int hw = H * W;
  const T* src = input + (size_t)nc * hw;
  if (kSmem) {
    for (int p = threadIdx.x; p < hw; p += kWslThreads) plane_s[p] = Elem<T>::to_f(src[p]);
    __syncthreads();
  }
  const int PP = PH * PW;
  const size_t set = (size_t)num_rois * channels * PP;
  const size_t set_h = (size_t)num_rois * PH, set_w = (size_t)num_rois * PW;
  const int total = (r1 - r0) * PP;
  for (int i = threadIdx.x; i < total; i += kWslThreads) {
    const int rl = i / PP;
    const int bin = i - rl * PP;
    const int ph = bin / PW, pw = bin - ph * PW;
    const int r = r0 + rl;
    if (ws.batch_idx[r] != n) continue;
    const LoopRects lr = rects[r];
    const size_t o = ((size_t)r * channels + c) * PP + bin;
    // pass 0: RoI rect -> RoI max and frame max (cells not strictly inside the inner rect)
    {
      const uint32_t hb = ws.hbins[(size_t)r * PH + ph], wb = ws.wbins[(size_t)r * PW + pw];
      float m = 0.f, mf = 0.f;  // maxval starts at 0 (ROILoopPool_cuda.cu:109-110)
      int mi = -1, mfi = -1;
      for (int h = range_lo(hb); h < range_hi(hb); ++h) {
        const bool in_h = h > lr.in0[1] && h < lr.in0[3];
        for (int w = range_lo(wb); w < range_hi(wb); ++w) {
          const int idx = h * W + w;
          const float v = plane_at<T, kSmem>(plane_s, src, idx);
          if (v > m) { m = v; mi = idx; }
          if (in_h && w > lr.in0[0] && w < lr.in0[2]) continue;
          if (v > mf) { mf = v; mfi = idx; }
        }
      }
      output[o] = Elem<T>::from_f(m);
      argmax[o] = mi;
      output[o + set] = Elem<T>::from_f(mf);
      argmax[o + set] = mfi;
    }
    // pass 1: outer rect -> context max (cells not strictly inside the RoI rect)
    {
      const uint32_t hb = ws.hbins[set_h + (size_t)r * PH + ph], wb = ws.wbins[set_w + (size_t)r * PW + pw];
      float mf = 0.f;
      int mfi = -1;
      for (int h = range_lo(hb); h < range_hi(hb); ++h) {
        const bool in_h = h > lr.in1[1] && h < lr.in1[3];
        for (int w = range_lo(wb); w < range_hi(wb); ++w) {
          if (in_h && w > lr.in1[0] && w < lr.in1[2]) continue;
          const int idx = h * W + w;
          const float v = plane_at<T, kSmem>(plane_s, src, idx);
          if (v > mf) { mf = v; mfi = idx; }
        }
      }
      output[o + 2 * set] = Elem<T>::from_f(mf);
      argmax[o + 2 * set] = mfi;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// MOIPool
// ---------------------------------------------------------------------------------------------
// mask[r, h, w] = 1 iff the cell lies in the RoI rectangle (inclusive ends) and some superpixel under the
// cell's footprint is switched on for this RoI (MOIPool_cuda.cu:151-214).
__global__ void __launch_bounds__(256)
moi_mask_kernel(const float* __restrict__ rois, const int* __restrict__ oh_labels, int len_labels,
                const int* __restrict__ superpixels, int Hs, int Ws, int batch, int H, int W, float scale,
                int num_rois, unsigned char* __restrict__ mask) {
  const int hw = H * W;
  const int r = blockIdx.y;
  const float* q = rois + (size_t)r * 5;
  const int b = (int)q[0];
  const RoiRect rc = roi_rect(q[1], q[2], q[3], q[4], scale);
  const float ss = (float)(1.0 * H / (double)Hs);  // MOIPool_cuda.cu:175
  const int* spp = superpixels + (size_t)b * Hs * Ws;
  const int* ohl = oh_labels + (size_t)r * len_labels;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < hw; p += gridDim.x * blockDim.x) {
    const int h = p / W, w = p - h * W;
    unsigned char v = 0;
    if (b >= 0 && b < batch && w >= rc.sw && w <= rc.ew && h >= rc.sh && h <= rc.eh) {
      int hh0 = (int)floorf(__fdiv_rn((float)h, ss)), ww0 = (int)floorf(__fdiv_rn((float)w, ss));
      int hh1 = (int)ceilf(__fdiv_rn((float)(h + 1), ss)), ww1 = (int)ceilf(__fdiv_rn((float)(w + 1), ss));
      hh0 = min(max(hh0, 0), Hs); hh1 = min(max(hh1, 0), Hs);
      ww0 = min(max(ww0, 0), Ws); ww1 = min(max(ww1, 0), Ws);
      int prev = -1;
      for (int hh = hh0; hh < hh1 && !v; ++hh)
        for (int ww = ww0; ww < ww1; ++ww) {
          const int id = __ldg(spp + (size_t)hh * Ws + ww);
          if (id == prev) continue;
          prev = id;
          if (id >= 0 && id < len_labels && __ldg(ohl + id) == 1) {
            v = 1;
            break;
          }
        }
    }
    mask[(size_t)r * hw + p] = v;
  }
}

template <typename T, bool kSmem>
__global__ void __launch_bounds__(kWslThreads)
moi_pool_fwd_kernel(const T* __restrict__ input, int channels, int H, int W, int PH, int PW, int splits,
                    PoolWorkspace ws, const unsigned char* __restrict__ mask,
                    const float* __restrict__ scales /* [2R]: mask scale, row scale; or NULL */, int num_rois,
                    T* __restrict__ output, int* __restrict__ argmax) {
  extern __shared__ float plane_s[];
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
  if (r0 >= r1) return;
  const int hw = H * W;
  const T* src = input + (size_t)nc * hw;
  if (kSmem) {
    for (int p = threadIdx.x; p < hw; p += kWslThreads) plane_s[p] = Elem<T>::to_f(src[p]);
    __syncthreads();
  }
  const int PP = PH * PW;
  const int total = (r1 - r0) * PP;
  for (int i = threadIdx.x; i < total; i += kWslThreads) {
    const int rl = i / PP;
    const int bin = i - rl * PP;
    const int ph = bin / PW, pw = bin - ph * PW;
    const int r = r0 + rl;
    if (ws.batch_idx[r] != n) continue;
    const uint32_t hb = ws.hbins[(size_t)r * PH + ph], wb = ws.wbins[(size_t)r * PW + pw];
    const unsigned char* mk = mask + (size_t)r * hw;
    float m = -FLT_MAX;
    int mi = -1;
    for (int h = range_lo(hb); h < range_hi(hb); ++h)
      for (int w = range_lo(wb); w < range_hi(wb); ++w) {
        const int idx = h * W + w;
        if (!__ldg(mk + idx)) continue;
        const float v = plane_at<T, kSmem>(plane_s, src, idx);
        if (v > m) { m = v; mi = idx; }
      }
    const size_t o = ((size_t)r * channels + c) * PP + bin;
    float v = mi == -1 ? 0.f : m;
    if (scales) v = __fmul_rn(__fmul_rn(v, scales[r]), scales[num_rois + r]);  // (f * mask_scale) * (objectness + 1)
    output[o] = Elem<T>::from_f(v);
    argmax[o] = mi;
  }
}

// ---- superpixel mask straight into row bit masks (visit-list path) ------------------------------------------------
// Which superpixels lie under a feature cell does not depend on the RoI: `moi_cell_ids_kernel` resolves it once per
// image (up to kMoiCellIds distinct ids per cell; more -> the cell is flagged and its footprint is rescanned per RoI).
// `moi_rowbits_kernel` then evaluates mask[r, h, w] = inside the RoI rectangle && some id of the cell is switched on
// (MOIPool_cuda.cu:151-214) for the rows of the RoI rectangle only, a warp per (RoI, row), and ballots the result
// into the 128-bit row word -- no [R, H, W] byte mask.
constexpr int kMoiCellIds = 8;

__device__ __forceinline__ void moi_footprint(int h, int w, float ss, int Hs, int Ws, int& hh0, int& hh1, int& ww0,
                                              int& ww1) {
  hh0 = (int)floorf(__fdiv_rn((float)h, ss));
  ww0 = (int)floorf(__fdiv_rn((float)w, ss));
  hh1 = (int)ceilf(__fdiv_rn((float)(h + 1), ss));
  ww1 = (int)ceilf(__fdiv_rn((float)(w + 1), ss));
  hh0 = min(max(hh0, 0), Hs); hh1 = min(max(hh1, 0), Hs);
  ww0 = min(max(ww0, 0), Ws); ww1 = min(max(ww1, 0), Ws);
}

__global__ void __launch_bounds__(256)
moi_cell_ids_kernel(const int* __restrict__ superpixels, int Hs, int Ws, int batch, int H, int W,
                    int* __restrict__ cell_ids /* [batch, H*W, kMoiCellIds] */,
                    unsigned char* __restrict__ cell_cnt /* [batch, H*W]; 255 = more than kMoiCellIds */) {
  const int hw = H * W;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (p >= hw) return;
  const int h = p / W, w = p - h * W;
  const float ss = (float)(1.0 * H / (double)Hs);  // MOIPool_cuda.cu:175
  int hh0, hh1, ww0, ww1;
  moi_footprint(h, w, ss, Hs, Ws, hh0, hh1, ww0, ww1);
  const int* spp = superpixels + (size_t)b * Hs * Ws;
  int ids[kMoiCellIds];
  int cnt = 0;
  bool over = false;
  for (int hh = hh0; hh < hh1 && !over; ++hh)
    for (int ww = ww0; ww < ww1; ++ww) {
      const int id = __ldg(spp + (size_t)hh * Ws + ww);
      bool seen = false;
#pragma unroll
      for (int k = 0; k < kMoiCellIds; ++k) seen |= (k < cnt && ids[k] == id);
      if (seen) continue;
      if (cnt == kMoiCellIds) { over = true; break; }
#pragma unroll
      for (int k = 0; k < kMoiCellIds; ++k)
        if (k == cnt) ids[k] = id;
      ++cnt;
    }
  int* dst = cell_ids + ((size_t)b * hw + p) * kMoiCellIds;
#pragma unroll
  for (int k = 0; k < kMoiCellIds; ++k) dst[k] = k < cnt ? ids[k] : -1;
  cell_cnt[(size_t)b * hw + p] = over ? 255 : (unsigned char)cnt;
}

__global__ void __launch_bounds__(256)
moi_rowbits_kernel(const float* __restrict__ rois, const int* __restrict__ oh_labels, int len_labels,
                   const int* __restrict__ superpixels, int Hs, int Ws, int batch, int H, int W, float scale,
                   int num_rois, const int* __restrict__ cell_ids, const unsigned char* __restrict__ cell_cnt,
                   uint4* __restrict__ rowbits) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= num_rois * H) return;
  const int r = row / H, h = row - r * H;
  const float* q = rois + (size_t)r * 5;
  const int b = (int)q[0];
  const RoiRect rc = roi_rect(q[1], q[2], q[3], q[4], scale);
  uint32_t bits[4] = {0u, 0u, 0u, 0u};
  if (b >= 0 && b < batch && h >= rc.sh && h <= rc.eh) {
    const int hw = H * W;
    const int* ohl = oh_labels + (size_t)r * len_labels;
    const float ss = (float)(1.0 * H / (double)Hs);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int w = k * 32 + lane;
      bool v = false;
      if (k * 32 <= rc.ew && k * 32 + 31 >= rc.sw) {  // warp-uniform: this 32-column chunk meets the rectangle
        if (w < W && w >= rc.sw && w <= rc.ew) {
          const size_t cell = (size_t)b * hw + (size_t)h * W + w;
          const int cnt = cell_cnt[cell];
          if (cnt != 255) {
            const int* ids = cell_ids + cell * kMoiCellIds;
            for (int j = 0; j < cnt && !v; ++j) {
              const int id = __ldg(ids + j);
              v = id >= 0 && id < len_labels && __ldg(ohl + id) == 1;
            }
          } else {  // many superpixels under this cell: scan the footprint like the reference
            int hh0, hh1, ww0, ww1;
            moi_footprint(h, w, ss, Hs, Ws, hh0, hh1, ww0, ww1);
            const int* spp = superpixels + (size_t)b * Hs * Ws;
            for (int hh = hh0; hh < hh1 && !v; ++hh)
              for (int ww = ww0; ww < ww1; ++ww) {
                const int id = __ldg(spp + (size_t)hh * Ws + ww);
                if (id >= 0 && id < len_labels && __ldg(ohl + id) == 1) { v = true; break; }
              }
          }
        }
        bits[k] = __ballot_sync(0xffffffffu, v);
      }
    }
  }
  if (lane == 0) rowbits[row] = make_uint4(bits[0], bits[1], bits[2], bits[3]);
}

// Row bit masks: rowbits[r, h] = 128-bit mask over the columns of row h (bit w = mask[r, h, w]); W <= 128.
// One warp per (RoI, row); four ballots.  15 MB of byte mask become 2.4 MB at config 2, and the pooling kernel
// reads one 16-byte word per bin row instead of one byte per cell.
__global__ void __launch_bounds__(256)
moi_pack_kernel(const unsigned char* __restrict__ mask, int rows /* R*H */, int W, uint4* __restrict__ rowbits) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const unsigned char* m = mask + (size_t)row * W;
  uint32_t b[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int w = k * 32 + lane;
    b[k] = __ballot_sync(0xffffffffu, w < W && m[w] != 0);
  }
  if (lane == 0) rowbits[row] = make_uint4(b[0], b[1], b[2], b[3]);
}

// MOIPool forward over row bit masks: only the switched-on cells of a bin are visited (row-major, strict >, so the
// value and the first arg-max are the reference's), no per-cell mask loads.
template <typename T>
__global__ void __launch_bounds__(kWslThreads)
moi_pool_fwd_bits_kernel(const T* __restrict__ input, int channels, int H, int W, int PH, int PW, int splits,
                         PoolWorkspace ws, const uint4* __restrict__ rowbits,
                         const float* __restrict__ scales, int num_rois, T* __restrict__ output,
                         int* __restrict__ argmax) {
  extern __shared__ float plane_s[];
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
  if (r0 >= r1) return;
  const int hw = H * W;
  const T* src = input + (size_t)nc * hw;
  for (int p = threadIdx.x; p < hw; p += kWslThreads) plane_s[p] = Elem<T>::to_f(src[p]);
  __syncthreads();
  const int PP = PH * PW;
  const int total = (r1 - r0) * PP;
  for (int i = threadIdx.x; i < total; i += kWslThreads) {
    const int rl = i / PP;
    const int bin = i - rl * PP;
    const int ph = bin / PW, pw = bin - ph * PW;
    const int r = r0 + rl;
    if (ws.batch_idx[r] != n) continue;
    const uint32_t hb = ws.hbins[(size_t)r * PH + ph], wb = ws.wbins[(size_t)r * PW + pw];
    const int w_lo = range_lo(wb), w_hi = range_hi(wb);
    const uint4* rm = rowbits + (size_t)r * H;
    float m = -FLT_MAX;
    int mi = -1;
    for (int h = range_lo(hb); h < range_hi(hb); ++h) {
      const uint4 q = __ldg(rm + h);
      for (int c0 = w_lo; c0 < w_hi; c0 += 32) {
        const int word = c0 >> 5, sh = c0 & 31;
        const uint32_t a = word == 0 ? q.x : word == 1 ? q.y : word == 2 ? q.z : q.w;
        const uint32_t b = word == 0 ? q.y : word == 1 ? q.z : word == 2 ? q.w : 0u;
        const int nb = min(32, w_hi - c0);
        uint32_t bits = __funnelshift_r(a, b, sh) & (nb >= 32 ? 0xffffffffu : ((1u << nb) - 1u));
        const int base = h * W + c0;
        while (bits) {
          const int j = __ffs(bits) - 1;
          bits &= bits - 1;
          const float v = plane_s[base + j];
          if (v > m) { m = v; mi = base + j; }
        }
      }
    }
    const size_t o = ((size_t)r * channels + c) * PP + bin;
    float v = mi == -1 ? 0.f : m;
    if (scales) v = __fmul_rn(__fmul_rn(v, scales[r]), scales[num_rois + r]);
    output[o] = Elem<T>::from_f(v);
    argmax[o] = mi;
  }
}

// ---- visit lists -----------------------------------------------------------------------------------------------
// Everything about "which cells does bin (r, ph, pw) look at" is channel independent, so it is resolved ONCE per call:
// per bin a packed list of the switched-on cells' row-major indices (uint16, row-major order, padded to a multiple of
// four with the sentinel index hw whose plane value is -FLT_MAX), and one descriptor word
//   desc[r * PP + bin] = (first quad of the list inside the RoI's slot) << 15 | number of quads   (both < 2^15).
// The per-plane kernel then runs one flat loop per bin: an 8-byte list load gives four indices, four shared-memory
// loads, four ordered compares -- no mask tests, no bit scans, no row loop.  A RoI's slot holds
// `cap_quads` uint2 entries (worst case: every cell of every bin switched on).
constexpr int kMoiListThreads = 128;
constexpr int kMoiSortMaxBins = 256;  // bins are sorted by list length up to 16 x 16 outputs per RoI

__device__ __forceinline__ uint32_t moi_row_bits(const uint4 q, int c0, int w_hi) {
  const int word = c0 >> 5, sh = c0 & 31;
  const uint32_t a = word == 0 ? q.x : word == 1 ? q.y : word == 2 ? q.z : q.w;
  const uint32_t b = word == 0 ? q.y : word == 1 ? q.z : word == 2 ? q.w : 0u;
  const int nb = min(32, w_hi - c0);
  return __funnelshift_r(a, b, sh) & (nb >= 32 ? 0xffffffffu : ((1u << nb) - 1u));
}

__global__ void __launch_bounds__(kMoiListThreads)
moi_list_kernel(int H, int W, int PH, int PW, PoolWorkspace ws, const uint4* __restrict__ rowbits,
                unsigned int cap_quads, uint32_t* __restrict__ desc, unsigned short* __restrict__ perm,
                uint2* __restrict__ lists, const float* __restrict__ row_scale, int use_mask_scale, int num_rois,
                float* __restrict__ scales /* [3R] or NULL: as moi_scale_kernel */,
                int range_max_mode /* ws.bdesc holds RoIPool descriptors: decide which bins keep them */,
                float* __restrict__ ms_out /* [R] or NULL: mask scale (1 when off) for the range-max kernel */,
                int* __restrict__ slow_total /* [1]: bins left to the visit lists (the gate) */) {
  __shared__ uint32_t s_quads[64 * 64 + 1];
  __shared__ unsigned char s_slow[64 * 64];
  const int r = blockIdx.x;
  const int PP = PH * PW;
  const uint4* rm = rowbits + (size_t)r * H;
  // 1. quads per bin
  for (int bin = threadIdx.x; bin < PP; bin += kMoiListThreads) {
    const int ph = bin / PW, pw = bin - ph * PW;
    const uint32_t hb = ws.hbins[(size_t)r * PH + ph], wb = ws.wbins[(size_t)r * PW + pw];
    int cnt = 0;
    for (int h = range_lo(hb); h < range_hi(hb); ++h) {
      const uint4 q = __ldg(rm + h);
      for (int c0 = range_lo(wb); c0 < range_hi(wb); c0 += 32) cnt += __popc(moi_row_bits(q, c0, range_hi(wb)));
    }
    s_quads[bin] = (uint32_t)(cnt + 3) >> 2;
    // range-max mode: a bin keeps its fast RoIPool descriptor only if every one of its cells is switched on; then the
    // range-max kernel answers it and the visit-list kernel skips it
    bool fast = false;
    if (range_max_mode) {
      const size_t id = (size_t)r * PP + bin;
      const int cells = (range_hi(hb) - range_lo(hb)) * (range_hi(wb) - range_lo(wb));
      fast = cnt > 0 && cnt == cells && (int)ws.bdesc[id] < 0;
      if (!fast) ws.bdesc[id] = 0u;  // the range-max main loop skips it
    }
    s_slow[bin] = !fast;
  }
  __syncthreads();
  // 2. exclusive scan (PP <= 4096; serial by one thread is ~50 adds at 7 x 7)
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    int valid = 0;  // bins with at least one switched-on cell
    int slow = 0;
    for (int bin = 0; bin < PP; ++bin) {
      const uint32_t n = s_quads[bin];
      s_quads[bin] = run;
      run += n;
      valid += n != 0;
      slow += s_slow[bin];
    }
    s_quads[PP] = run;
    if (range_max_mode) atomicAdd(slow_total, slow);
    if (ms_out) ms_out[r] = use_mask_scale ? __fmul_rn((float)PP, __frcp_rn((float)valid + 1.f)) : 1.f;
    if (scales) {  // roi_heads_jtsm.py:615-633, same arithmetic as moi_scale_kernel
      const float ms = use_mask_scale ? __fmul_rn((float)PP, __frcp_rn((float)valid + 1.f)) : 1.f;
      const float rs = row_scale ? row_scale[r] : 1.f;
      scales[r] = ms;
      scales[num_rois + r] = rs;
      scales[2 * (size_t)num_rois + r] = __fmul_rn(ms, rs);
    }
  }
  __syncthreads();
  // 3. fill the lists (slot layout in bin order) and emit descriptors in SORTED order: position k of a RoI holds the
  //    bin with the k-th most quads (ties: lower bin first), so the lanes of a warp run similar trip counts; the
  //    output run of a (RoI, channel) is still written by neighbouring lanes, only permuted.
  const int hw = H * W;
  uint2* slot = lists + (size_t)r * cap_quads;
  for (int bin = threadIdx.x; bin < PP; bin += kMoiListThreads) {
    const int ph = bin / PW, pw = bin - ph * PW;
    const uint32_t hb = ws.hbins[(size_t)r * PH + ph], wb = ws.wbins[(size_t)r * PW + pw];
    const uint32_t start = s_quads[bin];
    const uint32_t nq = s_quads[bin + 1] - start;
    // sort key: list length, the bins the range-max kernel answers last (key 0)
    const uint32_t key = s_slow[bin] ? nq + 1 : 0u;
    int pos = bin;
    if (PP <= kMoiSortMaxBins) {
      pos = 0;
      for (int o = 0; o < PP; ++o) {
        const uint32_t ko = s_slow[o] ? s_quads[o + 1] - s_quads[o] + 1 : 0u;
        pos += (ko > key) || (ko == key && o < bin);
      }
    }
    desc[(size_t)r * PP + pos] = (s_slow[bin] ? 0u : 0x80000000u) | (start << 15) | nq;  // bit 31: range-max bin
    perm[(size_t)r * PP + pos] = (unsigned short)bin;
    unsigned short* dst = (unsigned short*)(slot + start);
    int k = 0;
    for (int h = range_lo(hb); h < range_hi(hb); ++h) {
      const uint4 q = __ldg(rm + h);
      for (int c0 = range_lo(wb); c0 < range_hi(wb); c0 += 32) {
        uint32_t bits = moi_row_bits(q, c0, range_hi(wb));
        const int base = h * W + c0;
        while (bits) {
          dst[k++] = (unsigned short)(base + __ffs(bits) - 1);
          bits &= bits - 1;
        }
      }
    }
    for (; k < (int)(nq * 4); ++k) dst[k] = (unsigned short)hw;
  }
}

// NPL planes (channels of one image) per CTA: a list quad is loaded and unpacked once for NPL channels.
constexpr int kMoiFwdThreads = 512;

template <typename T, int NPL>
__global__ void __launch_bounds__(kMoiFwdThreads)
moi_pool_fwd_list_kernel(const T* __restrict__ input, int channels, int H, int W, int PP, int splits,
                         PoolWorkspace ws, const uint32_t* __restrict__ desc,
                         const unsigned short* __restrict__ perm, const uint2* __restrict__ lists,
                         unsigned int cap_quads, const float* __restrict__ scales, int num_rois,
                         const int* __restrict__ gate_count, int gate_total, int gate_pct,
                         int zero_floor /* ROILoopPool: the running maximum starts at 0 (ROILoopPool_cuda.cu:109) */,
                         T* __restrict__ output, int* __restrict__ argmax) {
  extern __shared__ float plane_s[];
  // range-max mode (device decision, same test in roi_pool_fwd_rmq_kernel): the fully switched-on bins (descriptor bit
  // 31) are answered there and skipped here; otherwise this kernel answers every bin
  const bool skip_fast = gate_count && moi_rmq_gate(gate_count, gate_total, gate_pct);
  const int item = blockIdx.x;
  const int s = item % splits;
  const int grp = item / splits;                 // group of NPL consecutive channels of one image
  const int groups_per_image = channels / NPL;   // channels % NPL == 0 (launcher)
  const int n = grp / groups_per_image;
  const int c0 = (grp - n * groups_per_image) * NPL;
  const int first = ws.first[n];
  const int last = -ws.neg_last[n];
  if (first > last) return;
  const int len = last - first + 1;
  const int r0 = first + (int)(((long long)len * s) / splits);
  const int r1 = first + (int)(((long long)len * (s + 1)) / splits);
  if (r0 >= r1) return;
  const int hw = H * W;
  const int pitch = hw + 1;  // + the padding cell
  const T* src = input + ((size_t)n * channels + c0) * hw;
  for (int p = threadIdx.x; p < NPL * hw; p += kMoiFwdThreads) {
    const int k = p / hw;
    plane_s[p + k] = Elem<T>::to_f(src[p]);
  }
  if (threadIdx.x < NPL) plane_s[threadIdx.x * pitch + hw] = -FLT_MAX;  // never beats the running maximum
  __syncthreads();
  const int total = (r1 - r0) * PP;
  for (int i = threadIdx.x; i < total; i += kMoiFwdThreads) {
    const int rl = i / PP;
    const int k = i - rl * PP;
    const int r = r0 + rl;
    if (ws.batch_idx[r] != n) continue;
    const uint32_t d = __ldg(desc + (size_t)r * PP + k);
    if (skip_fast && (int)d < 0) continue;
    const int bin = __ldg(perm + (size_t)r * PP + k);
    const uint2* lp = lists + (size_t)r * cap_quads + ((d >> 15) & 0xFFFFu);
    const int nq = (int)(d & 32767u);
    float m[NPL];
    int mi[NPL];
#pragma unroll
    for (int j = 0; j < NPL; ++j) m[j] = -FLT_MAX, mi[j] = -1;
    for (int q = 0; q < nq; ++q) {
      const uint2 e = __ldg(lp + q);
      const int i0 = e.x & 0xffff, i1 = e.x >> 16, i2 = e.y & 0xffff, i3 = e.y >> 16;
#pragma unroll
      for (int j = 0; j < NPL; ++j) {
        const float* pl = plane_s + j * pitch;
        const float v0 = pl[i0], v1 = pl[i1], v2 = pl[i2], v3 = pl[i3];
        if (v0 > m[j]) { m[j] = v0; mi[j] = i0; }
        if (v1 > m[j]) { m[j] = v1; mi[j] = i1; }
        if (v2 > m[j]) { m[j] = v2; mi[j] = i2; }
        if (v3 > m[j]) { m[j] = v3; mi[j] = i3; }
      }
    }
    float sc = 1.f, sr = 1.f;
    if (scales) sc = scales[r], sr = scales[num_rois + r];
#pragma unroll
    for (int j = 0; j < NPL; ++j) {
      const size_t o = ((size_t)r * channels + c0 + j) * PP + bin;
      if (zero_floor && !(m[j] > 0.f)) mi[j] = -1;
      float v = mi[j] == -1 ? 0.f : m[j];
      if (scales) v = __fmul_rn(__fmul_rn(v, sc), sr);  // (f * mask_scale) * (objectness + 1)
      output[o] = Elem<T>::from_f(v);
      argmax[o] = mi[j];
    }
  }
}
// Per-RoI scales of the JTSM heads (roi_heads_jtsm.py:615-633), fused into the pooling epilogue:
//   mask_scale = P*P * 1 / (#bins with at least one switched-on cell + 1)   -- the bins whose arg-max is not -1
//   row scale  = objectness_logits + 1 (caller supplied)
// scales[r] = mask scale (1 if off), scales[R + r] = row scale (1 if none), scales[2R + r] = their product
// (what the backward multiplies by).  One warp per RoI.
__global__ void moi_scale_kernel(int num_rois, int W, int hw, int PH, int PW, PoolWorkspace ws,
                                 const unsigned char* __restrict__ mask, const float* __restrict__ row_scale,
                                 int use_mask_scale, float* __restrict__ scales) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= num_rois) return;
  int valid = 0;
  if (use_mask_scale) {
    const unsigned char* mk = mask + (size_t)r * hw;
    for (int bin = lane; bin < PH * PW; bin += 32) {
      const int ph = bin / PW, pw = bin - ph * PW;
      const uint32_t hb = ws.hbins[(size_t)r * PH + ph], wb = ws.wbins[(size_t)r * PW + pw];
      bool any = false;
      for (int h = range_lo(hb); h < range_hi(hb) && !any; ++h)
        for (int w = range_lo(wb); w < range_hi(wb); ++w)
          if (mk[h * W + w]) {
            any = true;
            break;
          }
      valid += any;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) valid += __shfl_xor_sync(0xffffffffu, valid, o);
  }
  if (lane == 0) {
    const float ms = use_mask_scale ? __fmul_rn((float)(PH * PW), __frcp_rn((float)valid + 1.f)) : 1.f;
    const float rs = row_scale ? row_scale[r] : 1.f;
    scales[r] = ms;
    scales[num_rois + r] = rs;
    scales[2 * (size_t)num_rois + r] = __fmul_rn(ms, rs);
  }
}

}  // namespace jtsm

using namespace jtsm;

namespace {
int wsl_splits(int planes, int num_rois) {
  long long s = (4LL * num_sms() * 4 + planes - 1) / planes;
  const long long cap = num_rois / 32 > 1 ? num_rois / 32 : 1;
  s = s > cap ? cap : (s < 1 ? 1 : s);
  return (int)s;
}
}  // namespace

namespace jtsm {
__global__ void moi_prep_kernel(const float* __restrict__ rois, int num_rois, int batch, int H, int W,
                                float scale, int PH, int PW, PoolWorkspace ws) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= num_rois) return;
  const float* q = rois + (size_t)r * 5;
  const int b = (int)q[0];
  const RoiRect rc = roi_rect(q[1], q[2], q[3], q[4], scale);
  const int w_ = max(rc.ew - rc.sw + 1, 1), h_ = max(rc.eh - rc.sh + 1, 1);
  for (int p = 0; p < PH; ++p) ws.hbins[(size_t)r * PH + p] = bin_range(p, rc.sh, h_, PH, H);
  for (int p = 0; p < PW; ++p) ws.wbins[(size_t)r * PW + p] = bin_range(p, rc.sw, w_, PW, W);
  const bool ok = b >= 0 && b < batch;
  ws.batch_idx[r] = ok ? b : -1;
  if (ok) {
    atomicMin(ws.first + b, r);
    atomicMin(ws.neg_last + b, -r);
  }
}
}  // namespace jtsm

namespace {
// uint2 (= four uint16 cell indices) entries per RoI slot: the bins of a RoI overlap by at most one row / column,
// so their cells sum to at most (H + PH)(W + PW); + 3 padding entries per bin.  0 = visit lists not applicable.
size_t moi_cap_quads(int H, int W, int PH, int PW) {
  if (W > 128 || (long long)H * W > 65534) return 0;
  const size_t q = ((size_t)(H + PH) * (W + PW) + 3 * (size_t)PH * PW) / 4 + 1;
  return q < 32768 ? q : 0;
}
}  // namespace

extern "C" size_t jtsm_moi_pool_workspace_bytes(int batch, int height, int width, int num_rois, int pooled_h,
                                                int pooled_w) {
  if (batch < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0 || height < 0 || width < 0) return 0;
  return pool_workspace_bytes(batch, num_rois, pooled_h, pooled_w, 1) +
         align_up((size_t)num_rois * height * width, 256) + 256 +
         align_up((size_t)num_rois * height * sizeof(uint4), 256) +  // byte mask + row bit masks
         align_up((size_t)num_rois * pooled_h * pooled_w * sizeof(uint32_t), 256) +  // list descriptors
         align_up((size_t)num_rois * pooled_h * pooled_w * sizeof(unsigned short), 256) +  // sorted bin order
         align_up((size_t)num_rois * moi_cap_quads(height, width, pooled_h, pooled_w) * sizeof(uint2), 256) +
         align_up((size_t)batch * height * width * kMoiCellIds * sizeof(int), 256) +  // superpixel ids per cell
         align_up((size_t)batch * height * width, 256) +
         align_up((size_t)num_rois * sizeof(float), 256) + 256;  // mask scales + counter (range-max mode)
}

extern "C" int jtsm_moi_pool_forward(const void* input, const float* rois, const int* oh_labels, int len_labels,
                                     const int* superpixels, int height_sp, int width_sp, int dtype, int batch,
                                     int channels, int height, int width, int num_rois, float spatial_scale,
                                     int pooled_h, int pooled_w, const float* row_scale, int mask_scale,
                                     float* scale_out, void* output, int* argmax, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0 ||
      pooled_h > 64 || pooled_w > 64 || len_labels < 0 || height_sp <= 0 || width_sp <= 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (height > 32767 || width > 32767) return JTSM_ERR_UNSUPPORTED;
  if (num_rois == 0 || channels == 0) return JTSM_OK;
  if (!input || !rois || !oh_labels || !superpixels || !output || !argmax || !workspace || batch == 0 ||
      height == 0 || width == 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_moi_pool_workspace_bytes(batch, height, width, num_rois, pooled_h, pooled_w))
    return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  PoolWorkspace ws = carve_pool_workspace(workspace, batch, num_rois, pooled_h, pooled_w, 1);
  unsigned char* mask = (unsigned char*)(((uintptr_t)workspace + pool_workspace_bytes(batch, num_rois, pooled_h, pooled_w, 1) + 255) &
                                         ~(uintptr_t)255);
  const int hw = height * width;
  const int planes = batch * channels;
  const int splits = wsl_splits(planes, num_rois);
  const size_t smem = (size_t)hw * sizeof(float);
  const bool use_smem = smem <= 200 * 1024;
  const int algo = tuning("moi_algo", 0);  // 0 = auto (visit lists), 1 = row bit masks, 2 = byte-mask scan
  const bool use_bits = use_smem && width <= 128 && algo != 2;
  const size_t cap_quads = moi_cap_quads(height, width, pooled_h, pooled_w);
  const bool use_list = use_bits && cap_quads != 0 && algo == 0 && smem + sizeof(float) <= 200 * 1024;
  // range-max mode (roi_pool_rmq.cuh): fully switched-on bins are plain RoIPool bins and go to the range-max kernel,
  // the visit-list kernel answers the rest; decided per call on the device by the share of partially masked bins.
  // moi_rmq: -1 = off, 0..100 = use the range-max kernel when at most that many percent of the bins are left to the lists
  const int gate_pct = tuning("moi_rmq", 60);
  const bool use_rmq = use_list && gate_pct >= 0 &&
                       moi_rmq_eligible(batch, channels, height, width, num_rois, pooled_h, pooled_w);
  if (use_rmq) {
    const int rc = moi_rmq_prep(rois, num_rois, batch, channels, height, width, spatial_scale, pooled_h, pooled_w,
                                row_scale, ws, st);
    if (rc != JTSM_OK) return rc;
  } else {
    cudaMemsetAsync(ws.first, 0x7f, (size_t)batch * 2 * sizeof(int), st);
    moi_prep_kernel<<<(num_rois + 127) / 128, 128, 0, st>>>(rois, num_rois, batch, height, width, spatial_scale,
                                                            pooled_h, pooled_w, ws);
  }
  // workspace after the bin tables: byte mask | row bit masks | list descriptors | sorted bin order | lists |
  // per-cell superpixel ids | per-cell id counts
  const size_t npp = (size_t)num_rois * pooled_h * pooled_w;
  uint4* rowbits = (uint4*)(mask + align_up((size_t)num_rois * hw, 256));
  uint32_t* desc = (uint32_t*)((unsigned char*)rowbits + align_up((size_t)num_rois * height * sizeof(uint4), 256));
  unsigned short* perm = (unsigned short*)((unsigned char*)desc + align_up(npp * sizeof(uint32_t), 256));
  uint2* lists = (uint2*)((unsigned char*)perm + align_up(npp * sizeof(unsigned short), 256));
  int* cell_ids = (int*)((unsigned char*)lists + align_up((size_t)num_rois * cap_quads * sizeof(uint2), 256));
  unsigned char* cell_cnt = (unsigned char*)cell_ids + align_up((size_t)batch * hw * kMoiCellIds * sizeof(int), 256);
  float* ms_buf = (float*)(cell_cnt + align_up((size_t)batch * hw, 256));
  int* slow_total = (int*)((unsigned char*)ms_buf + align_up((size_t)num_rois * sizeof(float), 256));
  if ((row_scale || mask_scale) && !scale_out) return JTSM_ERR_INVALID_ARGUMENT;  // [3 * num_rois]
  const float* scales = (row_scale || mask_scale) ? scale_out : nullptr;
  if (use_list) {
    // RoI independent: superpixel ids under every cell; then the mask, straight into row bit masks
    moi_cell_ids_kernel<<<dim3((hw + 255) / 256, batch), 256, 0, st>>>(superpixels, height_sp, width_sp, batch, height,
                                                                      width, cell_ids, cell_cnt);
    const int rows = num_rois * height;
    moi_rowbits_kernel<<<(rows + 7) / 8, 256, 0, st>>>(rois, oh_labels, len_labels, superpixels, height_sp, width_sp,
                                                       batch, height, width, spatial_scale, num_rois, cell_ids,
                                                       cell_cnt, rowbits);
    if (use_rmq) cudaMemsetAsync(slow_total, 0, sizeof(int), st);
    moi_list_kernel<<<num_rois, kMoiListThreads, 0, st>>>(height, width, pooled_h, pooled_w, ws, rowbits,
                                                          (unsigned int)cap_quads, desc, perm, lists, row_scale,
                                                          mask_scale, num_rois, scales ? scale_out : nullptr,
                                                          use_rmq ? 1 : 0, use_rmq ? ms_buf : nullptr, slow_total);
    count_launch(use_rmq ? 3 : 4);  // (moi_rmq_prep counts its own launch)
    if (use_rmq) {
      MoiAux aux;
      aux.mask_scale = ms_buf;
      aux.slow_total = slow_total;
      aux.zero_floor = 0;
      aux.total_bins = (int)npp;
      aux.gate_pct = gate_pct;
      const int rc = moi_rmq_launch(dtype, input, batch, channels, height, width, num_rois, pooled_h, pooled_w, ws, aux,
                                    output, argmax, st);
      if (rc != JTSM_OK) return rc;
    }
  } else {
    dim3 mgrid((hw + 255) / 256 > 32 ? 32 : (hw + 255) / 256, num_rois);
    moi_mask_kernel<<<mgrid, 256, 0, st>>>(rois, oh_labels, len_labels, superpixels, height_sp, width_sp, batch,
                                           height, width, spatial_scale, num_rois, mask);
    count_launch(2);
    if (scales) {
      moi_scale_kernel<<<(num_rois + 7) / 8, 256, 0, st>>>(num_rois, width, hw, pooled_h, pooled_w, ws, mask,
                                                          row_scale, mask_scale, scale_out);
      count_launch();
    }
    if (use_bits) {
      const int rows = num_rois * height;
      moi_pack_kernel<<<(rows + 7) / 8, 256, 0, st>>>(mask, rows, width, rowbits);
      count_launch();
    }
  }
  JTSM_DISPATCH_DTYPE(dtype, T, {
    if (use_list) {
      const int npl_want = tuning("moi_planes", 2);
      const bool two = npl_want >= 2 && channels % 2 == 0 && 2 * (smem + sizeof(float)) <= 200 * 1024;
      const int npl = two ? 2 : 1;
      const size_t sm = npl * (smem + sizeof(float));
      const int groups = planes / npl;
      const int lsplits = wsl_splits(groups, num_rois);
      auto k = two ? moi_pool_fwd_list_kernel<T, 2> : moi_pool_fwd_list_kernel<T, 1>;
      if (sm > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
      k<<<groups * lsplits, kMoiFwdThreads, sm, st>>>((const T*)input, channels, height, width, pooled_h * pooled_w,
                                                      lsplits, ws, desc, perm, lists, (unsigned int)cap_quads, scales,
                                                      num_rois, use_rmq ? slow_total : nullptr, (int)npp, gate_pct, 0,
                                                      (T*)output, argmax);
    } else if (use_bits) {
      auto k = moi_pool_fwd_bits_kernel<T>;
      if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      k<<<planes * splits, kWslThreads, smem, st>>>((const T*)input, channels, height, width, pooled_h, pooled_w,
                                                     splits, ws, rowbits, scales, num_rois, (T*)output, argmax);
    } else if (use_smem) {
      auto k = moi_pool_fwd_kernel<T, true>;
      if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      k<<<planes * splits, kWslThreads, smem, st>>>((const T*)input, channels, height, width, pooled_h, pooled_w,
                                                     splits, ws, mask, scales, num_rois, (T*)output, argmax);
    } else {
      moi_pool_fwd_kernel<T, false><<<planes * splits, kWslThreads, 0, st>>>(
          (const T*)input, channels, height, width, pooled_h, pooled_w, splits, ws, mask, scales, num_rois,
          (T*)output, argmax);
    }
  });
  count_launch();
  return cuda_status();
}

// ---------------------------------------------------------------------------------------------
// ROILoopPool entry points (after the MOIPool machinery they reuse)
// ---------------------------------------------------------------------------------------------
namespace jtsm {
// ROILoopPool as a masked max-pool over 3R virtual RoIs (v = set * R + r, the op's own output order):
//   set 0: bins of the RoI rectangle, every cell allowed;
//   set 1: bins of the RoI rectangle, cells strictly inside the inner rectangle excluded (frame pool);
//   set 2: bins of the outer rectangle, cells strictly inside the RoI rectangle excluded (context pool)
// (ROILoopPool_cuda.cu:35-203).  One CTA per RoI writes the bin tables, the RoIPool range-max descriptors and the
// 128-bit row masks of its three virtual RoIs; moi_list_kernel / the range-max kernel / the visit-list kernel then run
// unchanged, with the running maximum floored at 0 (zero_floor).
__global__ void __launch_bounds__(128)
loop_virtual_prep_kernel(const float* __restrict__ rois, int num_rois, int batch, int H, int W, float scale,
                         float ratio, int PH, int PW, PoolWorkspace ws, uint4* __restrict__ rowbits) {
  __shared__ int s_rect[3][4];  // roi, inner, outer: (sw, sh, ew, eh)
  __shared__ int s_b;
  const int r = blockIdx.x;
  if (threadIdx.x == 0) {
    const float* q = rois + (size_t)r * 5;
    const float x1 = q[1], y1 = q[2], x2 = q[3], y2 = q[4];
    const float rw = __fsub_rn(x2, x1), rh = __fsub_rn(y2, y1);
    const float iw = __fdiv_rn(rw, ratio), ih = __fdiv_rn(rh, ratio);
    const float ow = __fmul_rn(rw, ratio), oh = __fmul_rn(rh, ratio);
    const float irw = __fsub_rn(rw, iw), irh = __fsub_rn(rh, ih);
    const float orw = __fsub_rn(ow, rw), orh = __fsub_rn(oh, rh);
    const float lim_w = (float)(1.0 * W / (double)scale), lim_h = (float)(1.0 * H / (double)scale);
    const RoiRect roi = roi_rect(x1, y1, x2, y2, scale);
    const RoiRect inner = roi_rect(clampf_(__fadd_rn(x1, __fdiv_rn(irw, 2.f)), 0.f, lim_w),
                                   clampf_(__fadd_rn(y1, __fdiv_rn(irh, 2.f)), 0.f, lim_h),
                                   clampf_(__fsub_rn(x2, __fdiv_rn(irw, 2.f)), 0.f, lim_w),
                                   clampf_(__fsub_rn(y2, __fdiv_rn(irh, 2.f)), 0.f, lim_h), scale);
    const RoiRect outer = roi_rect(clampf_(__fsub_rn(x1, __fdiv_rn(orw, 2.f)), 0.f, lim_w),
                                   clampf_(__fsub_rn(y1, __fdiv_rn(orh, 2.f)), 0.f, lim_h),
                                   clampf_(__fadd_rn(x2, __fdiv_rn(orw, 2.f)), 0.f, lim_w),
                                   clampf_(__fadd_rn(y2, __fdiv_rn(orh, 2.f)), 0.f, lim_h), scale);
    const RoiRect rc[3] = {roi, inner, outer};
    for (int k = 0; k < 3; ++k) {
      s_rect[k][0] = rc[k].sw; s_rect[k][1] = rc[k].sh; s_rect[k][2] = rc[k].ew; s_rect[k][3] = rc[k].eh;
    }
    const int b = (int)q[0];
    s_b = (b >= 0 && b < batch) ? b : -1;
    if (r == 0) *ws.ungrouped = batch > 1;  // virtual RoIs of an image are not one contiguous run
  }
  __syncthreads();
  const int hw = H * W, PP = PH * PW;
  for (int set = 0; set < 3; ++set) {
    const int v = set * num_rois + r;
    const int* base = s_rect[set == 2 ? 2 : 0];  // rectangle whose bins are pooled
    const int* ex = s_rect[set == 1 ? 1 : 0];    // exclusion rectangle (sets 1 and 2)
    const int w_ = max(base[2] - base[0] + 1, 1), h_ = max(base[3] - base[1] + 1, 1);
    for (int p = threadIdx.x; p < PH; p += 128) ws.hbins[(size_t)v * PH + p] = bin_range(p, base[1], h_, PH, H);
    for (int p = threadIdx.x; p < PW; p += 128) ws.wbins[(size_t)v * PW + p] = bin_range(p, base[0], w_, PW, W);
    if (threadIdx.x == 0) {
      ws.batch_idx[v] = s_b;
      ws.rscale[v] = 1.f;
      if (s_b >= 0) {
        atomicMin(ws.first + s_b, v);
        atomicMin(ws.neg_last + s_b, -v);
      }
    }
    for (int h = threadIdx.x; h < H; h += 128) {
      uint32_t m[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int nb = min(32, max(0, W - 32 * k));
        m[k] = nb >= 32 ? 0xffffffffu : ((1u << nb) - 1u);
      }
      if (set > 0 && h > ex[1] && h < ex[3]) {  // clear the columns strictly inside the exclusion rectangle
        const int lo = max(ex[0] + 1, 0), hi = min(ex[2] - 1, W - 1);
        for (int w = lo; w <= hi; ++w) m[w >> 5] &= ~(1u << (w & 31));
      }
      rowbits[(size_t)v * H + h] = make_uint4(m[0], m[1], m[2], m[3]);
    }
  }
  __syncthreads();  // the bin tables written above are read back below
  for (int j = threadIdx.x; j < 3 * PP; j += 128) {
    const int set = j / PP, bin = j - set * PP, ph = bin / PW, pw = bin - ph * PW;
    const int v = set * num_rois + r;
    ws.bdesc[(size_t)v * PP + bin] =
        rmq_bin_desc(ws.hbins[(size_t)v * PH + ph], ws.wbins[(size_t)v * PW + pw], W, hw);
  }
}
}  // namespace jtsm

namespace {
// Workspace of the virtual-RoI path: bin tables / descriptors for 3R virtual RoIs, their row bit masks, visit lists,
// sorted descriptors, unit mask scales and the gate counter.
struct LoopVirtLayout {
  size_t pool, rowbits, desc, perm, lists, ms, total;
  size_t cap_quads;
};
LoopVirtLayout loop_virt_layout(int batch, int height, int width, int num_rois, int PH, int PW) {
  LoopVirtLayout L;
  const size_t v = 3 * (size_t)num_rois, npp = v * PH * PW;
  L.cap_quads = moi_cap_quads(height, width, PH, PW);
  L.pool = pool_workspace_bytes(batch, (int)v, PH, PW, 1);
  L.rowbits = align_up(v * height * sizeof(uint4), 256);
  L.desc = align_up(npp * sizeof(uint32_t), 256);
  L.perm = align_up(npp * sizeof(unsigned short), 256);
  L.lists = align_up(v * L.cap_quads * sizeof(uint2), 256);
  L.ms = align_up(v * sizeof(float), 256);
  L.total = L.pool + L.rowbits + L.desc + L.perm + L.lists + L.ms + 512;
  return L;
}
bool loop_virt_ok(int batch, int channels, int height, int width, int num_rois, int PH, int PW) {
  return tuning("loop_algo", 0) == 0 && width <= 128 && moi_cap_quads(height, width, PH, PW) != 0 &&
         (long long)num_rois * 3 < 0x7FFFFFFFLL / (PH * PW) &&
         ((size_t)height * width + 1) * sizeof(float) * 2 <= 200 * 1024 &&
         moi_rmq_eligible(batch, channels, height, width, 3 * num_rois, PH, PW);
}
}  // namespace

extern "C" size_t jtsm_roi_loop_pool_workspace_bytes(int batch, int height, int width, int num_rois, int pooled_h,
                                                     int pooled_w) {
  if (batch < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0) return 0;
  const size_t direct = pool_workspace_bytes(batch, num_rois, pooled_h, pooled_w, 2) +
                        align_up((size_t)num_rois * sizeof(LoopRects), 256) + 256;
  const size_t virt = loop_virt_layout(batch, height, width, num_rois, pooled_h, pooled_w).total;
  return direct > virt ? direct : virt;
}

extern "C" int jtsm_roi_loop_pool_forward(const void* input, const float* rois, int dtype, int batch,
                                          int channels, int height, int width, int num_rois,
                                          float spatial_scale, float context_ratio, int pooled_h,
                                          int pooled_w, void* output, int* argmax, void* workspace,
                                          size_t workspace_bytes, void* stream) {
  if (batch < 0 || channels < 0 || height < 0 || width < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0 ||
      pooled_h > 64 || pooled_w > 64 || !(context_ratio > 0.f))
    return JTSM_ERR_INVALID_ARGUMENT;
  if (height > 32767 || width > 32767) return JTSM_ERR_UNSUPPORTED;
  if (num_rois == 0 || channels == 0) return JTSM_OK;
  if (!input || !rois || !output || !argmax || !workspace || batch == 0 || height == 0 || width == 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_roi_loop_pool_workspace_bytes(batch, height, width, num_rois, pooled_h, pooled_w))
    return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  if (loop_virt_ok(batch, channels, height, width, num_rois, pooled_h, pooled_w)) {
    // The three pools as 3R virtual RoIs of a masked max-pool (mask = cells outside the exclusion rectangle):
    // fully allowed bins -> RoIPool range-max kernel, the rest -> visit lists; values <= 0 give (0, -1).
    const int V = 3 * num_rois, PP = pooled_h * pooled_w, hw = height * width;
    const LoopVirtLayout L = loop_virt_layout(batch, height, width, num_rois, pooled_h, pooled_w);
    PoolWorkspace vw = carve_pool_workspace(workspace, batch, V, pooled_h, pooled_w, 1);
    char* p = (char*)(((uintptr_t)workspace + L.pool + 255) & ~(uintptr_t)255);
    uint4* rowbits = (uint4*)p; p += L.rowbits;
    uint32_t* desc = (uint32_t*)p; p += L.desc;
    unsigned short* perm = (unsigned short*)p; p += L.perm;
    uint2* lists = (uint2*)p; p += L.lists;
    float* ms_buf = (float*)p; p += L.ms;
    int* slow_total = (int*)p;
    cudaMemsetAsync(vw.first, 0x7f, (size_t)batch * 2 * sizeof(int), st);
    cudaMemsetAsync(vw.ungrouped, 0, 2 * sizeof(int), st);
    cudaMemsetAsync(slow_total, 0, sizeof(int), st);
    loop_virtual_prep_kernel<<<num_rois, 128, 0, st>>>(rois, num_rois, batch, height, width, spatial_scale,
                                                       context_ratio, pooled_h, pooled_w, vw, rowbits);
    moi_list_kernel<<<V, kMoiListThreads, 0, st>>>(height, width, pooled_h, pooled_w, vw, rowbits,
                                                   (unsigned int)L.cap_quads, desc, perm, lists, nullptr, 0, V, nullptr,
                                                   1, ms_buf, slow_total);
    count_launch(2);
    MoiAux aux;
    aux.mask_scale = ms_buf;
    aux.slow_total = slow_total;
    aux.total_bins = V * PP;
    aux.gate_pct = 100;  // the RoI set alone is a third of the bins and all of it is fully allowed
    aux.zero_floor = 1;
    int rc = moi_rmq_launch(dtype, input, batch, channels, height, width, V, pooled_h, pooled_w, vw, aux, output, argmax,
                            st);
    if (rc != JTSM_OK) return rc;
    const size_t plane_b = ((size_t)hw + 1) * sizeof(float);
    const bool two = channels % 2 == 0;
    const int npl = two ? 2 : 1;
    const int groups = batch * channels / npl;
    const int lsplits = wsl_splits(groups, V);
    JTSM_DISPATCH_DTYPE(dtype, T, {
      auto k = two ? moi_pool_fwd_list_kernel<T, 2> : moi_pool_fwd_list_kernel<T, 1>;
      if (npl * plane_b > 48 * 1024)
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(npl * plane_b));
      k<<<groups * lsplits, kMoiFwdThreads, npl * plane_b, st>>>(
          (const T*)input, channels, height, width, PP, lsplits, vw, desc, perm, lists, (unsigned int)L.cap_quads,
          nullptr, V, slow_total, V * PP, 100, 1, (T*)output, argmax);
    });
    count_launch();
    return cuda_status();
  }
  PoolWorkspace ws = carve_pool_workspace(workspace, batch, num_rois, pooled_h, pooled_w, 2);
  LoopRects* rects = (LoopRects*)(((uintptr_t)workspace + pool_workspace_bytes(batch, num_rois, pooled_h, pooled_w, 2) + 255) &
                                  ~(uintptr_t)255);
  cudaMemsetAsync(ws.first, 0x7f, (size_t)batch * 2 * sizeof(int), st);
  loop_pool_prep_kernel<<<(num_rois + 127) / 128, 128, 0, st>>>(rois, num_rois, batch, height, width, spatial_scale,
                                                                context_ratio, pooled_h, pooled_w, ws, rects);
  count_launch();
  const int planes = batch * channels;
  const int splits = wsl_splits(planes, num_rois);
  const size_t smem = (size_t)height * width * sizeof(float);
  const bool use_smem = smem <= 200 * 1024;
  JTSM_DISPATCH_DTYPE(dtype, T, {
    if (use_smem) {
      auto k = loop_pool_fwd_kernel<T, true>;
      if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      k<<<planes * splits, kWslThreads, smem, st>>>((const T*)input, channels, height, width, pooled_h, pooled_w,
                                                     splits, num_rois, ws, rects, (T*)output, argmax);
    } else {
      loop_pool_fwd_kernel<T, false><<<planes * splits, kWslThreads, 0, st>>>(
          (const T*)input, channels, height, width, pooled_h, pooled_w, splits, num_rois, ws, rects, (T*)output, argmax);
    }
  });
  count_launch();
  return cuda_status();
}

