// WSDDN two-stream MIL head math for sm_100a: per-image softmax(C, dim=1) * softmax(D, dim=0),
// proposal-sum image scores, clamp, binary cross-entropy, and the analytic backward.
//
// Reference (eager PyTorch composition, ~6 kernels per image + split/cat):
//   projects/WSL/wsl/modeling/roi_heads/fast_rcnn_wsddn.py:571-584  (scores)
//   projects/WSL/wsl/modeling/roi_heads/fast_rcnn_wsddn.py:364-379  (image scores, clamp)
//   projects/WSL/wsl/modeling/roi_heads/fast_rcnn_wsddn.py:346-362  (BCE mean | sum / N)
//
// Design: one CTA (1024 threads = 32 warps) per image.  Warp w owns rows w, w+32, ...;
// lane l owns class columns l, l+32, ... (KT register slots).  Row reductions are warp
// shuffles; column reductions are per-lane online (max, sum-exp) accumulators combined across
// the 32 warps through shared memory in a fixed order -> bit-reproducible run to run.
// The whole head is 2 launches (fwd+loss, bwd) instead of ~25.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace jtsm {

constexpr int kMilThreads = 1024;
constexpr int kMilWarps = kMilThreads / 32;

struct OnlineLse {  // running (max, sum exp(x - max))
  float m, s;
};
__device__ __forceinline__ void lse_push(OnlineLse& a, float x) {
  if (x > a.m) {
    a.s = a.s * expf(a.m - x) + 1.f;
    a.m = x;
  } else {
    a.s += expf(x - a.m);
  }
}
__device__ __forceinline__ void lse_merge(OnlineLse& a, const OnlineLse& b) {
  if (b.m == -INFINITY) return;
  if (a.m == -INFINITY) {
    a = b;
    return;
  }
  const float m = fmaxf(a.m, b.m);
  a.s = a.s * expf(a.m - m) + b.s * expf(b.m - m);
  a.m = m;
}

// Row softmax statistics of row `c` (this warp; lane holds KT values): returns max and 1/sum.
template <int KT>
__device__ __forceinline__ void row_stats(const float (&c)[KT], int K, int lane, float& rmax, float& rinv) {
  float m = -INFINITY;
#pragma unroll
  for (int t = 0; t < KT; ++t)
    if (lane + 32 * t < K) m = fmaxf(m, c[t]);
  m = warp_max(m);
  float s = 0.f;
#pragma unroll
  for (int t = 0; t < KT; ++t)
    if (lane + 32 * t < K) s += expf(c[t] - m);
  s = warp_sum(s);
  rmax = m;
  rinv = 1.f / s;
}

// torch.nn.functional.binary_cross_entropy element (log clamped at -100) and its derivative.
__device__ __forceinline__ float bce_elem(float y, float t) {
  const float ly = fmaxf(logf(y), -100.f);
  const float l1y = fmaxf(log1pf(-y), -100.f);
  return -(t * ly + (1.f - t) * l1y);
}
__device__ __forceinline__ float bce_grad(float y, float t) {
  return (y - t) / fmaxf((1.f - y) * y, 1e-12f);
}
// ---------------------------------------------------------------------------------------------
// A thread-block cluster of kMilCluster CTAs owns one image (a single CTA was issue bound:
// 94 us at R=2000, K=20 -- profiles/r1c).  CTA j takes rows r0 + j*32 + warp, stride 32*cluster.
// Column partials are combined warp -> CTA in shared memory, then CTA -> cluster through
// distributed shared memory, always in a fixed order (bit-reproducible run to run).
// Dynamic shared memory layout (floats): part[kMilWarps][K][2] | xchg[K][2] | fin[K][2]
// ---------------------------------------------------------------------------------------------
constexpr int kMilCluster = 8;

// Combine this CTA's per-warp (max, sumexp) partials and then the cluster's; result in fin[k][0..1].
__device__ __forceinline__ void cluster_combine_lse(cg::cluster_group& cluster, float* part, float* xchg, float* fin,
                                                    int K) {
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += kMilThreads) {
    OnlineLse a = {-INFINITY, 0.f};
    for (int w = 0; w < kMilWarps; ++w) {
      OnlineLse b = {part[(w * K + k) * 2 + 0], part[(w * K + k) * 2 + 1]};
      lse_merge(a, b);
    }
    xchg[k * 2 + 0] = a.m;
    xchg[k * 2 + 1] = a.s;
  }
  cluster.sync();
  const unsigned nr = cluster.num_blocks();
  for (int k = threadIdx.x; k < K; k += kMilThreads) {
    OnlineLse a = {-INFINITY, 0.f};
    for (unsigned j = 0; j < nr; ++j) {
      const float* rx = cluster.map_shared_rank(xchg, j);
      OnlineLse b = {rx[k * 2 + 0], rx[k * 2 + 1]};
      lse_merge(a, b);
    }
    fin[k * 2 + 0] = a.m;
    fin[k * 2 + 1] = a.s;
  }
  cluster.sync();  // everyone has read xchg before it is reused; fin visible CTA-wide
}

// Same for plain sums: part[kMilWarps][K] -> fin[k]
__device__ __forceinline__ void cluster_combine_sum(cg::cluster_group& cluster, float* part, float* xchg, float* fin,
                                                    int K) {
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += kMilThreads) {
    float t = 0.f;
    for (int w = 0; w < kMilWarps; ++w) t += part[w * K + k];
    xchg[k] = t;
  }
  cluster.sync();
  const unsigned nr = cluster.num_blocks();
  for (int k = threadIdx.x; k < K; k += kMilThreads) {
    float t = 0.f;
    for (unsigned j = 0; j < nr; ++j) t += cluster.map_shared_rank(xchg, j)[k];
    fin[k] = t;
  }
  cluster.sync();
}

template <int KT>
__global__ void __launch_bounds__(kMilThreads)
wsddn_fwd_kernel(const float* __restrict__ C, const float* __restrict__ D,
                 const int* __restrict__ seg, int num_images, int K,
                 const float* __restrict__ labels, int mean_loss, float* __restrict__ scores,
                 float* __restrict__ img_scores, float* __restrict__ col_stats,
                 float* __restrict__ loss, float* __restrict__ grad_img,
                 float* __restrict__ img_loss_partial, unsigned int* __restrict__ ticket) {
  extern __shared__ float sm[];
  cg::cluster_group cluster = cg::this_cluster();
  float* part = sm;
  float* xchg = sm + kMilWarps * K * 2;
  float* fin = xchg + K * 2;
  const int n = blockIdx.x / cluster.num_blocks();
  const int rank = cluster.block_rank();
  const int row_stride = kMilWarps * cluster.num_blocks();
  const int r0 = seg[n], r1 = seg[n + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rbeg = r0 + rank * kMilWarps + warp;

  // ---- pass A: column (max, sum-exp) of D over the image's rows
  OnlineLse col[KT];
#pragma unroll
  for (int t = 0; t < KT; ++t) col[t] = {-INFINITY, 0.f};
  for (int r = rbeg; r < r1; r += row_stride) {
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < K) lse_push(col[t], D[(size_t)r * K + k]);
    }
  }
#pragma unroll
  for (int t = 0; t < KT; ++t) {
    const int k = lane + 32 * t;
    if (k < K) {
      part[(warp * K + k) * 2 + 0] = col[t].m;
      part[(warp * K + k) * 2 + 1] = col[t].s;
    }
  }
  cluster_combine_lse(cluster, part, xchg, fin, K);
  float cmax[KT], cinv[KT];
#pragma unroll
  for (int t = 0; t < KT; ++t) {
    const int k = lane + 32 * t;
    cmax[t] = k < K ? fin[k * 2 + 0] : 0.f;
    const float cs = k < K ? fin[k * 2 + 1] : 0.f;
    cinv[t] = cs > 0.f ? 1.f / cs : 0.f;
    if (k < K && warp == 0 && rank == 0) {
      col_stats[((size_t)n * K + k) * 3 + 0] = cmax[t];
      col_stats[((size_t)n * K + k) * 3 + 1] = cs;
    }
  }

  // ---- pass B: scores = softmax_row(C) * softmax_col(D); column sums
  float ysum[KT];
#pragma unroll
  for (int t = 0; t < KT; ++t) ysum[t] = 0.f;
  for (int r = rbeg; r < r1; r += row_stride) {
    float c[KT], d[KT];
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      c[t] = k < K ? C[(size_t)r * K + k] : 0.f;
      d[t] = k < K ? D[(size_t)r * K + k] : 0.f;
    }
    float rmax, rinv;
    row_stats<KT>(c, K, lane, rmax, rinv);
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < K) {
        const float a = expf(c[t] - rmax) * rinv;
        const float b = expf(d[t] - cmax[t]) * cinv[t];
        const float s = a * b;
        scores[(size_t)r * K + k] = s;
        ysum[t] += s;
      }
    }
  }
  __syncthreads();  // fin (pass A) fully consumed before part is overwritten? part != fin; keeps phases tidy
#pragma unroll
  for (int t = 0; t < KT; ++t) {
    const int k = lane + 32 * t;
    if (k < K) part[warp * K + k] = ysum[t];
  }
  cluster_combine_sum(cluster, part, xchg, fin, K);
  if (rank != 0) return;
  float bce_part = 0.f;
  if (warp == 0) {
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < K) {
        const float y = fin[k];
        col_stats[((size_t)n * K + k) * 3 + 2] = y;
        const float yc = fminf(fmaxf(y, 1e-6f), 1.f - 1e-6f);
        img_scores[(size_t)n * K + k] = yc;
        if (labels) {
          const float tg = labels[(size_t)n * K + k];
          bce_part += bce_elem(yc, tg);
          // clamp passes gradient on [min, max] inclusive
          const float pass = (y >= 1e-6f && y <= 1.f - 1e-6f) ? 1.f : 0.f;
          const float norm = mean_loss ? 1.f / ((float)num_images * (float)K) : 1.f / (float)num_images;
          if (grad_img) grad_img[(size_t)n * K + k] = pass * bce_grad(yc, tg) * norm;
        }
      }
    }
    bce_part = warp_sum(bce_part);
    if (lane == 0 && labels && loss) {
      img_loss_partial[n] = bce_part;
      __threadfence();
      const unsigned int done = atomicAdd(ticket, 1u);
      if (done == (unsigned int)num_images - 1) {  // last image: fixed-order final sum
        __threadfence();
        float tot = 0.f;
        for (int i = 0; i < num_images; ++i) tot += ((volatile float*)img_loss_partial)[i];
        const float norm = mean_loss ? 1.f / ((float)num_images * (float)K) : 1.f / (float)num_images;
        *loss = tot * norm;
        *ticket = 0u;
      }
    }
  }
}

// Backward.  g_rk = grad_scores[r,k] (optional) + grad_img[n,k] * grad_loss (optional).
//   dC_rk = s_rk g_rk - a_rk * sum_j s_rj g_rj
//   dD_rk = s_rk g_rk - b_rk * sum_r' s_r'k g_r'k
template <int KT>
__global__ void __launch_bounds__(kMilThreads)
wsddn_bwd_kernel(const float* __restrict__ C, const float* __restrict__ D,
                 const float* __restrict__ scores, const float* __restrict__ col_stats,
                 const int* __restrict__ seg, int K, const float* __restrict__ grad_scores,
                 const float* __restrict__ grad_img, const float* __restrict__ grad_loss,
                 float* __restrict__ dC, float* __restrict__ dD) {
  extern __shared__ float sm[];
  cg::cluster_group cluster = cg::this_cluster();
  float* part = sm;
  float* xchg = sm + kMilWarps * K * 2;
  float* fin = xchg + K * 2;
  const int n = blockIdx.x / cluster.num_blocks();
  const int rank = cluster.block_rank();
  const int row_stride = kMilWarps * cluster.num_blocks();
  const int r0 = seg[n], r1 = seg[n + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rbeg = r0 + rank * kMilWarps + warp;
  const float gl = grad_loss ? *grad_loss : 1.f;
  float gimg[KT], cmax[KT], cinv[KT], colsg[KT];
#pragma unroll
  for (int t = 0; t < KT; ++t) {
    const int k = lane + 32 * t;
    gimg[t] = (grad_img && k < K) ? grad_img[(size_t)n * K + k] * gl : 0.f;
    cmax[t] = k < K ? col_stats[((size_t)n * K + k) * 3 + 0] : 0.f;
    const float cs = k < K ? col_stats[((size_t)n * K + k) * 3 + 1] : 0.f;
    cinv[t] = cs > 0.f ? 1.f / cs : 0.f;
    // sum_r s_rk * gimg_k = gimg_k * ysum_k
    colsg[t] = k < K ? gimg[t] * col_stats[((size_t)n * K + k) * 3 + 2] : 0.f;
  }
  if (grad_scores) {  // uniform branch: extra column reduction of s * grad_scores over the cluster
    float p[KT];
#pragma unroll
    for (int t = 0; t < KT; ++t) p[t] = 0.f;
    for (int r = rbeg; r < r1; r += row_stride) {
#pragma unroll
      for (int t = 0; t < KT; ++t) {
        const int k = lane + 32 * t;
        if (k < K) p[t] += scores[(size_t)r * K + k] * grad_scores[(size_t)r * K + k];
      }
    }
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < K) part[warp * K + k] = p[t];
    }
    cluster_combine_sum(cluster, part, xchg, fin, K);
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < K) colsg[t] += fin[k];
    }
  }
  for (int r = rbeg; r < r1; r += row_stride) {
    float c[KT];
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      c[t] = k < K ? C[(size_t)r * K + k] : 0.f;
    }
    float rmax, rinv;
    row_stats<KT>(c, K, lane, rmax, rinv);
    float sg[KT], rowsg = 0.f;
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      sg[t] = 0.f;
      if (k < K) {
        const float g = gimg[t] + (grad_scores ? grad_scores[(size_t)r * K + k] : 0.f);
        sg[t] = scores[(size_t)r * K + k] * g;
        rowsg += sg[t];
      }
    }
    rowsg = warp_sum(rowsg);
#pragma unroll
    for (int t = 0; t < KT; ++t) {
      const int k = lane + 32 * t;
      if (k < K) {
        const float a = expf(c[t] - rmax) * rinv;
        const float b = expf(D[(size_t)r * K + k] - cmax[t]) * cinv[t];
        dC[(size_t)r * K + k] = sg[t] - a * rowsg;
        dD[(size_t)r * K + k] = sg[t] - b * colsg[t];
      }
    }
  }
}

// cluster launch helper
template <typename Kern, typename... Args>
inline cudaError_t launch_cluster(Kern k, int num_images, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(num_images * kMilCluster));
  cfg.blockDim = dim3(kMilThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kMilCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, k, args...);
}

}  // namespace jtsm

using namespace jtsm;

extern "C" size_t jtsm_wsddn_workspace_bytes(int num_images) {
  return num_images < 0 ? 0 : align_up((size_t)num_images * 4 + 16, 256) + 256;
}

extern "C" int jtsm_wsddn_forward(const float* cls_logits, const float* det_logits,
                                  const int* seg_offsets, int num_images, int num_rois,
                                  int num_classes, const float* labels_onehot, int mean_loss,
                                  float* scores, float* img_scores, float* col_stats, float* loss,
                                  float* grad_img, void* workspace, size_t workspace_bytes,
                                  void* stream) {
  if (num_images < 0 || num_rois < 0 || num_classes <= 0 || num_classes > 256)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (num_images == 0) return JTSM_OK;
  if (!cls_logits && num_rois > 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (!seg_offsets || !scores || !img_scores || !col_stats || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_wsddn_workspace_bytes(num_images)) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  char* p = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  unsigned int* ticket = (unsigned int*)p;
  float* partial = (float*)(p + 16);
  cudaMemsetAsync(ticket, 0, 4, st);
  const size_t smem = ((size_t)kMilWarps * num_classes * 2 + 4 * num_classes) * sizeof(float);
  JTSM_DISPATCH_KT(num_classes, KT, {
    auto k = wsddn_fwd_kernel<KT>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    launch_cluster(k, num_images, smem, st, cls_logits, det_logits, seg_offsets, num_images, num_classes,
                   labels_onehot, mean_loss, scores, img_scores, col_stats, loss, grad_img, partial, ticket);
  });
  count_launch();
  return cuda_status();
}

extern "C" int jtsm_wsddn_backward(const float* cls_logits, const float* det_logits,
                                   const float* scores, const float* col_stats,
                                   const int* seg_offsets, int num_images, int num_rois,
                                   int num_classes, const float* grad_scores, const float* grad_img,
                                   const float* grad_loss, float* grad_cls, float* grad_det,
                                   void* stream) {
  if (num_images < 0 || num_rois < 0 || num_classes <= 0 || num_classes > 256)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (num_images == 0 || num_rois == 0) return JTSM_OK;
  if (!cls_logits || !det_logits || !scores || !col_stats || !seg_offsets || !grad_cls || !grad_det)
    return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = ((size_t)kMilWarps * num_classes * 2 + 4 * num_classes) * sizeof(float);
  JTSM_DISPATCH_KT(num_classes, KT, {
    auto k = wsddn_bwd_kernel<KT>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    launch_cluster(k, num_images, smem, st, cls_logits, det_logits, scores, col_stats, seg_offsets, num_classes,
                   grad_scores, grad_img, grad_loss, grad_cls, grad_det);
  });
  count_launch();
  return cuda_status();
}
