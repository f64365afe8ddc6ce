// Per-class (batched) greedy NMS, fully on device, for sm_100a.
//
// Reference: detectron2/layers/nms.py:10-30 (batched_nms: per-class torchvision.ops.nms, keep
// indices returned in descending-score order) and torchvision.ops.nms (greedy, suppress
// IoU > thr, candidates visited in stable descending-score order).
// Callers: fast_rcnn_inference_single_image (projects/WSL/wsl/modeling/roi_heads/fast_rcnn_oicr.py:95-163),
// get_pgt_mist (roi_heads_oicr.py:563-568).
//
// torchvision's GPU path builds an n x n/64 bitmask, copies it to the host and scans it on the
// CPU.  Here: two stable radix sorts give (a) the global descending-score order and (b) the
// per-class segments in that order; one CTA per class then runs greedy NMS over its segment in
// 64-box blocks (intra-block resolve by one warp on a 64x64 bitmask, kept boxes of the block
// suppress the tail in parallel); a flagged select emits the keep list.  No host sync.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>

#include <cooperative_groups.h>

#include "common.cuh"

namespace jtsm {

namespace cg = cooperative_groups;

constexpr int kNmsThreads = 512;
constexpr int kNmsCluster = 8;  // CTAs per class segment: the tail suppression is split over them by 64-box block

__device__ __forceinline__ unsigned int desc_key(float s) {
  // ascending radix order of this key == descending score order
  unsigned int u = __float_as_uint(s);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending-orderable
  return ~u;
}

constexpr unsigned int kPadKey = 0xFFFFFFFFu;  // padding entries (index >= *num_valid): last in score order, own class

__global__ void nms_keys_kernel(const float* __restrict__ scores, int n, const int* __restrict__ num_valid,
                                unsigned int* __restrict__ keys, int* __restrict__ vals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const int nv = num_valid ? min(*num_valid, n) : n;
    keys[i] = i < nv ? desc_key(scores[i]) : kPadKey;
    vals[i] = i;
  }
}

// cls1[pos] = class of the pos-th best box; pos1[pos] = pos
__global__ void nms_gather_class_kernel(const long long* __restrict__ idxs, const int* __restrict__ order1,
                                        int n, const int* __restrict__ num_valid, unsigned int* __restrict__ cls1,
                                        int* __restrict__ pos1) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const int nv = num_valid ? min(*num_valid, n) : n;
    const int src = order1[i];
    cls1[i] = src >= nv ? kPadKey : (idxs ? (unsigned int)idxs[src] : 0u);
    pos1[i] = i;
  }
}

// After the class sort: gather boxes into (class, score) order and find segment starts.
__global__ void nms_segments_kernel(const float* __restrict__ boxes, const int* __restrict__ order1,
                                    const int* __restrict__ order2, const unsigned int* __restrict__ cls2,
                                    int n, float4* __restrict__ sboxes, int* __restrict__ seg_start,
                                    int* __restrict__ seg_end, int* __restrict__ num_seg) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n) return;
  const unsigned int c = cls2[q];
  if (c == kPadKey) return;  // padding: no box, no segment
  sboxes[q] = *(const float4*)(boxes + (size_t)order1[order2[q]] * 4);
  if (q == 0 || c != cls2[q - 1]) {
    int lo = q + 1, hi = n;  // first position after q whose class differs (classes are sorted)
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (cls2[mid] == c) lo = mid + 1; else hi = mid;
    }
    const int slot = atomicAdd(num_seg, 1);
    seg_start[slot] = q;
    seg_end[slot] = lo;
  }
}

// torchvision nms IoU test (CPU kernel op order): inter / (area_i + area_j - inter) > thr
__device__ __forceinline__ bool suppresses(const float4 a, float area_a, const float4 b, float thr) {
  const float w = fmaxf(0.f, __fsub_rn(fminf(a.z, b.z), fmaxf(a.x, b.x)));
  const float h = fmaxf(0.f, __fsub_rn(fminf(a.w, b.w), fmaxf(a.y, b.y)));
  const float inter = __fmul_rn(w, h);
  const float area_b = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
  const float uni = __fsub_rn(__fadd_rn(area_a, area_b), inter);
  // inter / uni > thr  <=>  inter > thr * uni for uni > 0; the product and the quotient are each within one rounding
  // of their real values, so a relative margin of 1e-5 decides all but the closest calls without the division
  // (the exact quotient decides those, and every degenerate case: uni <= 0, non-finite, thr <= 0)
  const float t = __fmul_rn(thr, uni);
  if (uni > 1e-20f && uni < 1e30f && t > 0.f) {
    if (inter > __fmul_rn(t, 1.00001f)) return true;
    if (inter < __fmul_rn(t, 0.99999f)) return false;
  }
  return __fdiv_rn(inter, uni) > thr;
}

// CLUSTERS of kNmsCluster CTAs stride over class segments [seg_start, seg_end) (any order).
// Per 64-box block of a segment, every CTA of the cluster redundantly (A) stages the block's boxes in shared memory and
// computes the 64 x 64 suppression bits, (B) resolves the block with warp 0 -- the 64 mask rows are read up front
// (independent of the running `dead` word, so the loads pipeline), the dependent chain is one AND / OR per box -- and
// compacts the kept boxes; then (C) the kept boxes suppress the tail, split over the cluster: tail block t belongs to
// CTA t % kNmsCluster, which owns its removed bits.  The removed bits of the block being resolved are read from their
// owner through distributed shared memory; one cluster barrier per block orders (C) before the next (B).
// (Round 1: one CTA per class, a single thread resolved the block with a global-memory load per kept box.)
__global__ void __cluster_dims__(kNmsCluster, 1, 1) __launch_bounds__(kNmsThreads)
nms_segment_kernel(const float4* __restrict__ sboxes, const int* __restrict__ seg_start,
                   const int* __restrict__ seg_end, const int* __restrict__ num_seg,
                   const int* __restrict__ order2, int n, float thr,
                   unsigned char* __restrict__ keep_flag /* by global score rank */) {
  extern __shared__ unsigned int s_removed[];  // (n + 31) / 32 words (+ 2: a block's pair of words is always readable)
  __shared__ unsigned long long s_mask[64];
  __shared__ float4 s_box[64];
  __shared__ float4 s_kbox[64];
  __shared__ float s_karea[64];
  __shared__ unsigned long long s_kept;
  cg::cluster_group cluster = cg::this_cluster();
  const int cr = (int)cluster.block_rank();
  const int cid = blockIdx.x / kNmsCluster, ncl = gridDim.x / kNmsCluster;
  const int ns = *num_seg;
  for (int seg = cid; seg < ns; seg += ncl) {
    const int s = seg_start[seg], m = seg_end[seg] - s;
    const float4* B = sboxes + s;
    for (int i = threadIdx.x; i < (m + 31) / 32 + 2; i += kNmsThreads) s_removed[i] = 0u;
    cluster.sync();  // every CTA's bits are cleared before anyone reads them (and nobody still reads the last segment's)
    const int nblk = (m + 63) / 64;
    for (int blk = 0; blk < nblk; ++blk) {
      const int base = blk * 64;
      const int cnt = min(64, m - base);
      if (threadIdx.x < cnt) s_box[threadIdx.x] = B[base + threadIdx.x];
      __syncthreads();
      // A: 64x64 suppression bits inside the block (thread = (row i, 8-column group))
      {
        const int i = threadIdx.x >> 3, part = threadIdx.x & 7;
        unsigned long long bits = 0ull;
        if (i < cnt) {
          const float4 a = s_box[i];
          const float area_a = __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(a.w, a.y));
          for (int jj = 0; jj < 8; ++jj) {
            const int j = part * 8 + jj;
            if (j > i && j < cnt && suppresses(a, area_a, s_box[j], thr)) bits |= 1ull << j;
          }
        }
        // combine the 8 partial masks of row i (8 consecutive lanes); executed by every lane
        bits |= __shfl_xor_sync(0xffffffffu, bits, 1);
        bits |= __shfl_xor_sync(0xffffffffu, bits, 2);
        bits |= __shfl_xor_sync(0xffffffffu, bits, 4);
        if (part == 0 && i < 64) s_mask[i] = bits;
      }
      __syncthreads();
      // B: resolve by warp 0 (all lanes run the same chain); the block's removed bits live in CTA blk % kNmsCluster
      if (threadIdx.x < 32) {
        const unsigned int* owner = cluster.map_shared_rank(s_removed, blk % kNmsCluster);
        unsigned long long dead = (unsigned long long)owner[base >> 5] | ((unsigned long long)owner[(base >> 5) + 1] << 32);
        if (cnt < 64) dead |= ~0ull << cnt;
        unsigned long long kept = 0ull;
#pragma unroll 16
        for (int i = 0; i < 64; ++i) {
          const unsigned long long row = s_mask[i];  // rows >= cnt are zero
          const bool alive = !((dead >> i) & 1ull);
          kept |= alive ? (1ull << i) : 0ull;
          dead |= alive ? row : 0ull;
        }
        // the kept boxes, compacted in block order; their keep flags (written by one CTA of the cluster)
        for (int i = threadIdx.x; i < cnt; i += 32) {
          if ((kept >> i) & 1ull) {
            const int pos = __popcll(kept & ((1ull << i) - 1ull));
            const float4 a = s_box[i];
            s_kbox[pos] = a;
            s_karea[pos] = __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(a.w, a.y));
            if (cr == 0) keep_flag[order2[s + base + i]] = 1;
          }
        }
        if (threadIdx.x == 0) s_kept = kept;
      }
      __syncthreads();
      // C: kept boxes of this block suppress this CTA's share of the tail (blocks t > blk with t % cluster == rank)
      const int nk = __popcll(s_kept);
      int t0 = blk + 1 + ((cr - (blk + 1)) % kNmsCluster + kNmsCluster) % kNmsCluster;  // first owned block after blk
      for (int t = t0 + (threadIdx.x >> 6) * kNmsCluster; t < nblk; t += (kNmsThreads >> 6) * kNmsCluster) {
        const int j = t * 64 + (threadIdx.x & 63);
        if (j >= m || ((s_removed[j >> 5] >> (j & 31)) & 1u)) continue;
        const float4 b = B[j];
        bool kill = false;
        for (int k = 0; k < nk && !kill; ++k) kill = suppresses(s_kbox[k], s_karea[k], b, thr);
        if (kill) atomicOr(&s_removed[j >> 5], 1u << (j & 31));
      }
      cluster.sync();  // (C) of every CTA is complete before the next block is resolved
    }
  }
}

__global__ void nms_finish_kernel(const int* __restrict__ sel, const int* __restrict__ num_sel,
                                  long long* __restrict__ keep, int n, long long* __restrict__ num_keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = *num_sel;
  if (i < n) keep[i] = i < k ? (long long)sel[i] : -1;
  if (i == 0) *num_keep = k;
}

struct NmsWs {
  unsigned int *keys_a, *keys_b, *cls_a, *cls_b;
  int *vals_a, *order1, *pos_a, *order2, *seg_start, *seg_end, *num_seg, *sel, *num_sel;
  float4* sboxes;
  unsigned char* keep_flag;
  void* cub_temp;
  size_t cub_bytes;
};

inline size_t nms_cub_bound(int n) { return (size_t)n * 16 + (4u << 20); }

inline size_t nms_layout(int n, void* base, NmsWs* w) {
  char* p = (char*)(((uintptr_t)base + 255) & ~(uintptr_t)255);
  char* p0 = p;
  auto take = [&](size_t bytes) {
    char* q = p;
    p += align_up(bytes, 256);
    return q;
  };
  const size_t n4 = (size_t)n * 4;
  NmsWs t;
  t.keys_a = (unsigned int*)take(n4);
  t.keys_b = (unsigned int*)take(n4);
  t.cls_a = (unsigned int*)take(n4);
  t.cls_b = (unsigned int*)take(n4);
  t.vals_a = (int*)take(n4);
  t.order1 = (int*)take(n4);
  t.pos_a = (int*)take(n4);
  t.order2 = (int*)take(n4);
  t.seg_start = (int*)take(n4);
  t.seg_end = (int*)take(n4);
  t.sel = (int*)take(n4);
  t.sboxes = (float4*)take((size_t)n * 16);
  // zero-initialised block: keep flags, counters
  t.keep_flag = (unsigned char*)take((size_t)n);
  t.num_seg = (int*)take(16);
  t.num_sel = t.num_seg + 1;
  t.cub_temp = take(nms_cub_bound(n));
  t.cub_bytes = nms_cub_bound(n);
  if (w) *w = t;
  return (size_t)(p - p0) + 256;
}

}  // namespace jtsm

using namespace jtsm;

extern "C" size_t jtsm_nms_workspace_bytes(int num_boxes) {
  if (num_boxes < 0) return 0;
  return nms_layout(num_boxes, nullptr, nullptr);
}

namespace jtsm {
// The whole NMS pipeline.  num_valid (DEVICE int, optional): only boxes [0, *num_valid) take part -- for callers whose
// candidate count is produced on the device (fast_rcnn_inference: threshold + compaction); n is then the capacity.
int nms_run(const float* boxes, const float* scores, const long long* idxs, int n, const int* num_valid,
            float iou_threshold, long long* keep, long long* num_keep, void* workspace, size_t workspace_bytes,
            cudaStream_t st) {
  if (workspace_bytes < jtsm_nms_workspace_bytes(n)) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  NmsWs w;
  nms_layout(n, workspace, &w);
  const int tb = 256, gb = (n + tb - 1) / tb;
  // zero the contiguous scratch block [keep_flag .. num_sel]
  cudaMemsetAsync(w.keep_flag, 0, (size_t)((char*)w.cub_temp - (char*)w.keep_flag), st);
  nms_keys_kernel<<<gb, tb, 0, st>>>(scores, n, num_valid, w.keys_a, w.vals_a);
  size_t need = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, need, w.keys_a, w.keys_b, w.vals_a, w.order1, n, 0, 32, st);
  if (need > w.cub_bytes) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cub::DeviceRadixSort::SortPairs(w.cub_temp, need, w.keys_a, w.keys_b, w.vals_a, w.order1, n, 0, 32, st);
  nms_gather_class_kernel<<<gb, tb, 0, st>>>(idxs, w.order1, n, num_valid, w.cls_a, w.pos_a);
  need = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, need, w.cls_a, w.cls_b, w.pos_a, w.order2, n, 0, 32, st);
  if (need > w.cub_bytes) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cub::DeviceRadixSort::SortPairs(w.cub_temp, need, w.cls_a, w.cls_b, w.pos_a, w.order2, n, 0, 32, st);
  nms_segments_kernel<<<gb, tb, 0, st>>>(boxes, w.order1, w.order2, w.cls_b, n, w.sboxes, w.seg_start, w.seg_end,
                                         w.num_seg);
  // clusters stride over segments; the segment count is only known on the device
  int clusters = 2 * num_sms() / kNmsCluster;
  if (clusters > n) clusters = n;
  const int max_seg = (clusters < 1 ? 1 : clusters) * kNmsCluster;
  const size_t smem = ((size_t)n / 32 + 4) * 4;
  if (smem > 160 * 1024) return JTSM_ERR_UNSUPPORTED;
  if (smem > 40 * 1024)
    cudaFuncSetAttribute(nms_segment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  nms_segment_kernel<<<max_seg, kNmsThreads, smem, st>>>(w.sboxes, w.seg_start, w.seg_end, w.num_seg, w.order2, n,
                                                         iou_threshold, w.keep_flag);
  need = 0;
  cub::DeviceSelect::Flagged(nullptr, need, w.order1, w.keep_flag, w.sel, w.num_sel, n, st);
  if (need > w.cub_bytes) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cub::DeviceSelect::Flagged(w.cub_temp, need, w.order1, w.keep_flag, w.sel, w.num_sel, n, st);
  nms_finish_kernel<<<gb, tb, 0, st>>>(w.sel, w.num_sel, keep, n, num_keep);
  count_launch(6);
  return cuda_status();
}
}  // namespace jtsm

extern "C" int jtsm_batched_nms(const float* boxes, const float* scores, const long long* idxs,
                                int num_boxes, float iou_threshold, long long* keep,
                                long long* num_keep, void* workspace, size_t workspace_bytes,
                                void* stream) {
  if (num_boxes < 0 || !num_keep) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  if (num_boxes == 0) {
    cudaMemsetAsync(num_keep, 0, sizeof(long long), st);
    return cuda_status();
  }
  if (!boxes || !scores || !keep || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  return nms_run(boxes, scores, idxs, num_boxes, nullptr, iou_threshold, keep, num_keep, workspace, workspace_bytes, st);
}
