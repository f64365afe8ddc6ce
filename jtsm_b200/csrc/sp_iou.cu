// Superpixel-set IoU + Matcher for JTSM's mask-branch label assignment.
//
// Reference: pairwise_iou_sp, projects/WSL/wsl/modeling/roi_heads/roi_heads_jtsm.py:49-73 -- materialises three
// [n1, n2, S] integer tensors (n1 pseudo-gt targets, n2 proposals, S superpixels) to compute
//   inter[i, j] = sum_s oh1[i, s] * oh2[j, s] * cnt[s],  area1[i] = sum_s oh1[i, s] * cnt[s],  area2[j] likewise,
//   iou = float(inter) / float(area1 + area2 - inter)
// and feeds it to Matcher (detectron2/modeling/matcher.py:61-103) in label_and_sample_proposals_wsl (:1627-1634).
// Here: one launch, HBM bound on the single read of oh2 (n2 x S x 4 bytes); a warp owns a proposal row, the
// targets' weighted rows (oh1 * cnt) sit in shared memory, sums are exact 64-bit integers, the conversions and the
// division are round-to-nearest like the CPU's, so IoU values, matched index and label are bit-exact.
#include "common.cuh"

namespace jtsm {

constexpr int kSpTile = 8;  // targets per pass

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// torch.max(dim=0) order: first maximum; NaN beats everything (first NaN wins)
__device__ __forceinline__ bool better(float v, float best) { return v > best || (isnan(v) && !isnan(best)); }

__global__ void __launch_bounds__(256) pairwise_iou_sp_kernel(
    const int* __restrict__ oh1, int n1, const int* __restrict__ oh2, int n2, const long long* __restrict__ cnt,
    int S, float* __restrict__ iou, float t0, float t1, int l0, int l1, int l2, int num_thr,
    long long* __restrict__ matched_idx, signed char* __restrict__ matched_label) {
  extern __shared__ long long sp_smem[];
  long long* s_cnt = sp_smem;              // [S]
  long long* s_w1 = s_cnt + S;             // [kSpTile][S]  oh1 * cnt
  __shared__ long long s_area1[kSpTile];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int s = threadIdx.x; s < S; s += blockDim.x) s_cnt[s] = cnt[s];

  // per-warp running match over the target tiles (rows handled: warp-strided, same set in every tile pass)
  const int rows_per_cta = (n2 + gridDim.x - 1) / gridDim.x;
  const int r_begin = blockIdx.x * rows_per_cta, r_end = min(n2, r_begin + rows_per_cta);

  for (int g0 = 0; g0 < n1 || g0 == 0; g0 += kSpTile) {
    const int gt = min(kSpTile, n1 - g0);
    __syncthreads();
    for (int i = threadIdx.x; i < gt * S; i += blockDim.x) {
      const int g = i / S, s = i - g * S;
      s_w1[g * S + s] = (long long)oh1[(size_t)(g0 + g) * S + s] * s_cnt[s];
    }
    __syncthreads();
    if (warp < gt) {
      long long a = 0;
      for (int s = lane; s < S; s += 32) a += s_w1[warp * S + s];
      a = warp_sum_ll(a);
      if (lane == 0) s_area1[warp] = a;
    }
    __syncthreads();
    for (int r = r_begin + warp; r < r_end; r += nwarp) {
      const int* row = oh2 + (size_t)r * S;
      long long inter[kSpTile];
#pragma unroll
      for (int g = 0; g < kSpTile; ++g) inter[g] = 0;
      long long area2 = 0;
      for (int s = lane; s < S; s += 32) {
        const long long v = (long long)__ldg(row + s);
        area2 += v * s_cnt[s];
#pragma unroll
        for (int g = 0; g < kSpTile; ++g)
          if (g < gt) inter[g] += v * s_w1[g * S + s];
      }
      area2 = warp_sum_ll(area2);
      float best = 0.f;
      int best_g = -1;
      if (g0 > 0 && matched_idx) {  // carry the running best of the earlier tiles (kept in the outputs)
        best_g = (int)matched_idx[r];
        best = iou[(size_t)best_g * n2 + r];
      }
#pragma unroll
      for (int g = 0; g < kSpTile; ++g) {
        if (g >= gt) break;
        const long long in = warp_sum_ll(inter[g]);
        const float q = __fdiv_rn(__ll2float_rn(in), __ll2float_rn(s_area1[g] + area2 - in));
        if (lane == 0) iou[(size_t)(g0 + g) * n2 + r] = q;
        if (best_g < 0 || better(q, best)) best = q, best_g = g0 + g;
      }
      if (lane == 0 && matched_idx) {
        matched_idx[r] = n1 > 0 ? best_g : 0;
        if (g0 + kSpTile >= n1 && matched_label) {
          // Matcher: labels start at 1; interval i = [thr[i-1], thr[i]) with -inf / +inf at the ends
          int lab = 1;
          if (n1 == 0) {
            lab = l0;
          } else if (num_thr == 1) {
            if (best < t0) lab = l0; else if (best >= t0) lab = l1;
          } else {
            if (best < t0) lab = l0; else if (best >= t0 && best < t1) lab = l1; else if (best >= t1) lab = l2;
          }
          matched_label[r] = (signed char)lab;
        }
      }
    }
    if (n1 == 0) break;
  }
}

}  // namespace jtsm

using namespace jtsm;

extern "C" int jtsm_pairwise_iou_sp(const int* oh1, int n1, const int* oh2, int n2, const long long* cnt_sp, int S,
                                    float* iou, float iou_t0, float iou_t1, int label0, int label1, int label2,
                                    int num_thresholds, long long* matched_idx, signed char* matched_label,
                                    void* stream) {
  if (n1 < 0 || n2 < 0 || S <= 0 || num_thresholds < 1 || num_thresholds > 2) return JTSM_ERR_INVALID_ARGUMENT;
  if (n2 == 0) return JTSM_OK;
  if (!oh2 || !cnt_sp || (n1 > 0 && (!oh1 || !iou))) return JTSM_ERR_INVALID_ARGUMENT;
  if (matched_label && !matched_idx) return JTSM_ERR_INVALID_ARGUMENT;
  const size_t smem = (size_t)(1 + kSpTile) * S * sizeof(long long);
  if (smem > 200 * 1024) return JTSM_ERR_INVALID_ARGUMENT;  // S <= 2844
  cudaFuncSetAttribute(pairwise_iou_sp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int grid = std::min(num_sms() * 2, (n2 + 7) / 8);
  pairwise_iou_sp_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(oh1, n1, oh2, n2, cnt_sp, S, iou, iou_t0, iou_t1,
                                                                   label0, label1, label2, num_thresholds,
                                                                   matched_idx, matched_label);
  count_launch();
  return cuda_status();
}
