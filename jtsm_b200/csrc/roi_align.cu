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
#include "common.cuh"

namespace jtsm {

struct AlignGeom {  // ROIAlign_cpu.cpp:139-166
  float start_h, start_w, bin_h, bin_w;
  int grid_h, grid_w, batch, pad;
};

struct AlignWorkspace {
  AlignGeom* geom;  // [R]
  int* first;       // [batch]
  int* neg_last;    // [batch]
  int* ungrouped;   // [1]
};

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

constexpr int kAlignThreads = 512;

// grid: persistent CTAs over (plane, chunk) items.  kSmem: plane staged in shared memory.
template <typename T, bool kSmem>
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
      for (int p = threadIdx.x; p < hw; p += kAlignThreads) plane_s[p] = Elem<T>::to_f(src[p]);
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
      for (int iy = 0; iy < g.grid_h; ++iy) {
        const float y = sample_coord(g.start_h, ph, g.bin_h, iy, g.grid_h);
        for (int ix = 0; ix < g.grid_w; ++ix) {
          const float x = sample_coord(g.start_w, pw, g.bin_w, ix, g.grid_w);
          Taps t;
          float val = 0.f;
          if (bilinear_taps(H, W, y, x, t)) {
            float v1, v2, v3, v4;
            if (kSmem) {
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

template <typename T>
__global__ void align_convert_from_f32_kernel(const float* __restrict__ src, T* __restrict__ dst, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = Elem<T>::from_f(src[i]);
}

}  // namespace jtsm

using namespace jtsm;

namespace {
int align_prep(const float* rois, int num_rois, int batch, float scale, int PH, int PW, int sampling_ratio,
               int aligned, const AlignWorkspace& ws, cudaStream_t st) {
  cudaMemsetAsync(ws.first, 0x7f, (size_t)batch * 2 * sizeof(int), st);
  cudaMemsetAsync(ws.ungrouped, 0, sizeof(int), st);
  roi_align_prep_kernel<<<(num_rois + 127) / 128, 128, 0, st>>>(rois, num_rois, batch, scale, PH, PW,
                                                                sampling_ratio, aligned, ws);
  count_launch();
  return cuda_status();
}

template <typename T>
int launch_align_fwd(const void* input, int batch, int channels, int H, int W, int num_rois, int PH, int PW,
                     const AlignWorkspace& ws, void* output, cudaStream_t st) {
  const int planes = batch * channels;
  const size_t smem = (size_t)H * W * sizeof(float);
  const bool use_smem = smem <= 200 * 1024 && !tuning("roi_align_force_gmem", 0);
  int per_sm = use_smem ? (int)((220 * 1024) / (smem + 1024)) : 4;
  per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
  const int max_grid = num_sms() * per_sm;
  int chunks = (8 * max_grid + planes - 1) / planes;
  const int cap = num_rois / 32 > 1 ? num_rois / 32 : 1;
  chunks = chunks > cap ? cap : (chunks < 1 ? 1 : chunks);
  const int items = planes * chunks;
  const int grid = items < max_grid ? items : max_grid;
  if (use_smem) {
    auto k = roi_align_fwd_kernel<T, true>;
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

extern "C" size_t jtsm_roi_align_workspace_bytes(int dtype, int batch, int channels, int height, int width,
                                                 int num_rois, int backward) {
  if (batch < 0 || num_rois < 0) return 0;
  size_t b = align_workspace_bytes(batch, num_rois);
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
  int rc = align_prep(rois, num_rois, batch, spatial_scale, pooled_h, pooled_w, sampling_ratio, aligned, ws, st);
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
    const int cpp = channels * pooled_h * pooled_w;
    int ysplit = (cpp + 255) / 256;
    const int want = (4 * num_sms() * 8 + num_rois - 1) / num_rois;
    if (ysplit > want) ysplit = want < 1 ? 1 : want;
    dim3 grid(num_rois, ysplit);
    JTSM_DISPATCH_DTYPE(dtype, T, {
      roi_align_bwd_kernel<T><<<grid, 256, 0, st>>>((const T*)grad_output, stride_n, stride_c, stride_h, stride_w,
                                                    channels, height, width, pooled_h, pooled_w, num_rois, ws, acc);
    });
    count_launch();
    rc = cuda_status();
    if (rc != JTSM_OK) return rc;
  }
  if (dtype == JTSM_BF16) {
    align_convert_from_f32_kernel<__nv_bfloat16><<<kNumSMs * 8, 256, 0, st>>>(acc, (__nv_bfloat16*)grad_input, n_in);
    count_launch();
  } else if (dtype == JTSM_F16) {
    align_convert_from_f32_kernel<__half><<<kNumSMs * 8, 256, 0, st>>>(acc, (__half*)grad_input, n_in);
    count_launch();
  }
  return cuda_status();
}
