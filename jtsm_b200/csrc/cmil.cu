// ROIMerge and ROILabel on the device (the CMIL / UWSOD heads' two native ops).
//
// Reference: projects/WSL/wsl/layers/csrc/ROIMerge/ROIMerge_cpu.cpp:33-287 and ROILabel/ROILabel_cpu.cpp:16-194 -- both
// CPU only (ROIMerge.h:48-53, ROILabel.h:29-34 raise for CUDA tensors), so the reference's heads copy the scores and the
// R x R IoU matrix to the host on every iteration (fast_rcnn_wsddn.py:666, roi_heads_cmil.py:419-431).
//
// ROIMerge: rank the RoIs by objectness, grow greedy cliques among the top 200 inside windows of 40 ranks (a candidate
//   joins when its IoU with every member so far is >= lambda), every other RoI is its own cluster; the cluster's C / D
//   rows are the means of its members'.  Here: CUB radix sort (stable, descending) -> ONE CTA builds the 200 x 200
//   ">= lambda" bit matrix in shared memory and warp 0 runs the greedy pass with ballots (a round per ACCEPTED member: the
//   lowest still-compatible candidate is final because the member set only grows) -> ids of the rest by a block scan in
//   index order -> a kernel sums each cluster's rows in ascending RoI order, each term divided first, exactly the
//   reference's running sum: ids, counts and merged scores are bit-identical.
// ROILabel: per switched-on class the top_k best not-yet-chosen rows are seeds; every RoI takes the seed of highest IoU
//   (first maximum), becomes foreground (>= fg) or background ([lo, hi)) while the per-call caps last, in a visiting order
//   the reference draws with std::random_shuffle(time(0)).  Here the visiting order is an input permutation (caller's RNG)
//   and the two caps are exclusive prefix counts over that order: one CTA, bit-identical labels and weights.
#include <cub/cub.cuh>

#include "common.cuh"

namespace jtsm {
namespace {

constexpr int kMergeMaxTop = 256;     // top_k <= 256 (reference: 200)
constexpr int kMergeMaxWindow = 64;   // window <= 64 (reference: 40)
constexpr int kMergeWords = kMergeMaxTop / 32;
constexpr int kLabelMaxSeeds = 256;

__global__ void iota_kernel(int* __restrict__ v, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) v[i] = i;
}

struct MergeWs {
  float* keys_out;
  int* vals_in;
  int* order;    // [R] RoI indices, descending score
  int* members;  // [kMergeMaxTop * kMergeMaxWindow] members of the top clusters, ascending RoI index
  int* owner;    // [R] the RoI of a singleton cluster id
  void* cub_tmp;
  size_t cub_bytes;
};

size_t merge_cub_bytes(int n) {
  size_t b = 0;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, b, (const float*)nullptr, (float*)nullptr, (const int*)nullptr,
                                            (int*)nullptr, n);
  return b;
}

MergeWs carve_merge_ws(void* ws, int n) {
  char* p = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  MergeWs w;
  const size_t rb = align_up((size_t)n * 4, 256);
  w.keys_out = (float*)p; p += rb;
  w.vals_in = (int*)p; p += rb;
  w.order = (int*)p; p += rb;
  w.owner = (int*)p; p += rb;
  w.members = (int*)p; p += align_up((size_t)kMergeMaxTop * kMergeMaxWindow * 4, 256);
  w.cub_tmp = p;
  w.cub_bytes = merge_cub_bytes(n);
  return w;
}

// One CTA of 1024 threads.  stats: [0] num_id, [1] num_top_id, [2] max_clique, [3] min_clique.
__global__ void __launch_bounds__(1024) roi_merge_clique_kernel(const float* __restrict__ J, const int* __restrict__ order,
                                                                int R, float lambda, int top_k, int window,
                                                                int* __restrict__ I, int* __restrict__ IC,
                                                                int* __restrict__ members, int* __restrict__ owner,
                                                                int* __restrict__ stats) {
  __shared__ int s_ord[kMergeMaxTop];
  __shared__ int s_id[kMergeMaxTop];
  __shared__ int s_seed_rank[kMergeMaxTop];  // rank of the first member of top cluster id
  __shared__ int s_cnt[kMergeMaxTop];
  __shared__ unsigned s_adj[kMergeMaxTop][kMergeWords];
  __shared__ int s_num_top;
  __shared__ int s_scan[32];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int T = min(top_k, R);
  for (int a = tid; a < T; a += blockDim.x) {
    s_ord[a] = order[a];
    s_id[a] = -1;
    s_cnt[a] = 0;
  }
  __syncthreads();
  // bit (a, b) = !(J[ord[a], ord[b]] < lambda): candidate a against member b (ROIMerge_cpu.cpp:131)
  for (int w = tid; w < T * kMergeWords; w += blockDim.x) {
    const int a = w / kMergeWords, b0 = (w % kMergeWords) * 32;
    unsigned bits = 0;
    const float* row = J + (size_t)s_ord[a] * R;
    for (int k = 0; k < 32 && b0 + k < T; k++)
      if (!(row[s_ord[b0 + k]] < lambda)) bits |= 1u << k;
    s_adj[a][w % kMergeWords] = bits;
  }
  __syncthreads();
  if (warp == 0) {
    int cur = 0;
    for (int t = 0; t < T; t++) {
      if (s_id[t] != -1) continue;  // warp-uniform (shared memory, written under __syncwarp below)
      const int end = min(t + window, T);
      // members as a bit set over ranks [t, t + 64): lane l watches candidates t + l and t + 32 + l
      unsigned long long M = 1ull;
      int c0 = t + lane, c1 = t + 32 + lane;
      bool open0 = c0 < end && lane != 0 && s_id[c0] == -1;
      bool open1 = c1 < end && s_id[c1] == -1;
      // the 64-bit slice [t, t + 64) of a candidate's adjacency row
      auto slice = [&](int a) -> unsigned long long {
        const int w0 = t >> 5, sh = t & 31;
        unsigned long long lo = s_adj[a][w0];
        unsigned long long mid = w0 + 1 < kMergeWords ? s_adj[a][w0 + 1] : 0u;
        unsigned long long hi = w0 + 2 < kMergeWords ? s_adj[a][w0 + 2] : 0u;
        unsigned long long v = (lo | (mid << 32)) >> sh;
        if (sh) v |= hi << (64 - sh);
        return v;
      };
      const unsigned long long r0 = open0 ? slice(c0) : 0ull, r1 = open1 ? slice(c1) : 0ull;
      if (lane == 0) s_id[t] = cur;
      int count = 1;
      while (true) {
        // a candidate that fails against the current members fails for good; the lowest one that passes is accepted
        open0 = open0 && ((M & ~r0) == 0ull);
        open1 = open1 && ((M & ~r1) == 0ull);
        const unsigned b0 = __ballot_sync(0xffffffffu, open0), b1 = __ballot_sync(0xffffffffu, open1);
        if (!(b0 | b1)) break;
        const int pick = b0 ? __ffs(b0) - 1 : 32 + __ffs(b1) - 1;  // offset from t
        M |= 1ull << pick;
        if (pick < 32) { if (lane == pick) { open0 = false; s_id[t + pick] = cur; } }
        else if (lane == pick - 32) { open1 = false; s_id[t + pick] = cur; }
        count++;
      }
      if (lane == 0) { s_seed_rank[cur] = t; s_cnt[cur] = count; }
      cur++;
      __syncwarp();
    }
    if (lane == 0) s_num_top = cur;
  }
  __syncthreads();
  const int num_top = s_num_top;
  // ids / counts of the top clusters; member lists in ascending RoI index (the order of the reference's running sums)
  for (int a = tid; a < T; a += blockDim.x) I[s_ord[a]] = s_id[a];
  for (int id = warp; id < num_top; id += blockDim.x >> 5) {
    const int t = s_seed_rank[id];
    const int a0 = t + lane, a1 = t + 32 + lane;
    const int n0 = (a0 < T && s_id[a0] == id) ? s_ord[a0] : 0x7fffffff;
    const int n1 = (a1 < T && a1 < t + window && s_id[a1] == id) ? s_ord[a1] : 0x7fffffff;
    int p0 = 0, p1 = 0;
    for (int l = 0; l < 32; l++) {
      const int m0 = __shfl_sync(0xffffffffu, n0, l), m1 = __shfl_sync(0xffffffffu, n1, l);
      p0 += (m0 < n0) + (m1 < n0);
      p1 += (m0 < n1) + (m1 < n1);
    }
    if (n0 != 0x7fffffff) members[id * kMergeMaxWindow + p0] = n0;
    if (n1 != 0x7fffffff) members[id * kMergeMaxWindow + p1] = n1;
    if (lane == 0) IC[id] = s_cnt[id];
  }
  if (tid == 0) s_carry = 0;
  __syncthreads();  // I of the top RoIs is visible to the whole CTA (same block; global writes + barrier)
  // everything else: its own cluster, numbered in index order after the top clusters (:149-158)
  for (int base = 0; base < R; base += blockDim.x) {
    const int n = base + tid;
    const int single = (n < R && I[n] == -1) ? 1 : 0;
    int incl = single;
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int v = s_scan[lane];
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
      }
      s_scan[lane] = v;
    }
    __syncthreads();
    const int before = s_carry + (warp ? s_scan[warp - 1] : 0) + incl - single;
    if (single) {
      const int id = num_top + before;
      I[n] = id;
      IC[id] = 1;
      owner[id] = n;
    }
    __syncthreads();
    if (tid == 0) s_carry += s_scan[31];
    __syncthreads();
  }
  const int num_id = num_top + s_carry;
  for (int id = num_id + tid; id < R; id += blockDim.x) IC[id] = 0;
  if (tid == 0) {
    int mx = 0, mn = T;  // :183-193
    for (int i = 0; i < num_top; i++) {
      mx = max(mx, s_cnt[i]);
      mn = min(mn, s_cnt[i]);
    }
    stats[0] = num_id;
    stats[1] = num_top;
    stats[2] = mx;
    stats[3] = mn;
  }
}

// MC[id] = running sum over the members in ascending RoI index of C[n] / count (ROIMerge_cpu.cpp:198-205).
__global__ void roi_merge_sum_kernel(const float* __restrict__ C, const float* __restrict__ D, int R, int K,
                                     const int* __restrict__ IC, const int* __restrict__ members,
                                     const int* __restrict__ owner, const int* __restrict__ stats,
                                     float* __restrict__ MC, float* __restrict__ MD) {
  const int num_id = stats[0], num_top = stats[1];
  const long long total = (long long)R * K;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int id = (int)(e / K), c = (int)(e % K);
    float mc = 0.f, md = 0.f;
    if (id < num_top) {
      const int cnt = IC[id];
      const float d = (float)cnt;
      for (int j = 0; j < cnt; j++) {
        const int n = members[id * kMergeMaxWindow + j];
        mc += __fdiv_rn(C[(size_t)n * K + c], d);
        md += __fdiv_rn(D[(size_t)n * K + c], d);
      }
    } else if (id < num_id) {
      const int n = owner[id];
      mc = 0.f + C[(size_t)n * K + c];  // 0 + x / 1 (the addition turns -0 into +0 like the reference's)
      md = 0.f + D[(size_t)n * K + c];
    }
    MC[e] = mc;
    MD[e] = md;
  }
}

__global__ void roi_merge_bwd_kernel(const float* __restrict__ GMC, const float* __restrict__ GMD,
                                     const int* __restrict__ I, const int* __restrict__ IC, int R, int K,
                                     float* __restrict__ GC, float* __restrict__ GD) {
  const long long total = (long long)R * K;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(e / K), c = (int)(e % K);
    const int id = I[n];
    const float d = (float)IC[id];
    GC[e] = __fdiv_rn(GMC[(size_t)id * K + c], d);
    GD[e] = __fdiv_rn(GMD[(size_t)id * K + c], d);
  }
}


// ------------------------------------------------------------------------------------------------ ROILabel
// (value desc, index asc) block arg-max helper: every thread passes its best; result broadcast from shared memory.
struct BestRow {
  float v;
  int n;
};
__device__ __forceinline__ BestRow better(BestRow a, BestRow b) {
  // larger value wins; equal values: the lower row (the reference's strict '<' scan keeps the first maximum)
  if (b.n >= 0 && (a.n < 0 || b.v > a.v || (b.v == a.v && b.n < a.n))) return b;
  return a;
}

// One CTA of 1024 threads.  stats float[4]: fg_rois, bg_rois, fg_weight, bg_weight of this call.
__global__ void __launch_bounds__(1024) roi_label_kernel(const float* __restrict__ S, int Ks, const float* __restrict__ U,
                                                         const float* __restrict__ L, const float* __restrict__ CW,
                                                         const int* __restrict__ perm, int R, int K, float fg, float bg_hi,
                                                         float bg_lo, int num_pos, int num_neg, int top_k,
                                                         int* __restrict__ RL, float* __restrict__ RW,
                                                         float* __restrict__ stats, unsigned char* __restrict__ elig,
                                                         int* __restrict__ cls_tmp) {
  __shared__ int s_seed_n[kLabelMaxSeeds];
  __shared__ int s_seed_c[kLabelMaxSeeds];
  __shared__ float s_seed_p[kLabelMaxSeeds];
  __shared__ int s_num_seeds;
  __shared__ BestRow s_best[32];
  __shared__ int s_scan[2][32];
  __shared__ int s_carry[2];
  __shared__ float s_wsum[2][32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_num_seeds = 0;
  __syncthreads();
  // seeds (ROILabel_cpu.cpp:84-106): classes ascending, top_k each, rows already chosen (by any class) excluded
  for (int c = 0; c < K; c++) {
    if (L[c] != 1.f) continue;  // uniform
    for (int k = 0; k < top_k; k++) {
      const int ns = s_num_seeds;
      if (ns >= kLabelMaxSeeds) break;
      BestRow b{-FLT_MAX, -1};
      for (int n = tid; n < R; n += blockDim.x) {
        const float v = S[(size_t)n * Ks + c];
        if (!(b.n < 0 ? -FLT_MAX < v : b.v < v)) continue;
        bool taken = false;
        for (int i = 0; i < ns; i++) taken |= s_seed_n[i] == n;
        if (!taken) b = BestRow{v, n};
      }
      for (int o = 16; o; o >>= 1) {
        BestRow x{__shfl_xor_sync(0xffffffffu, b.v, o), __shfl_xor_sync(0xffffffffu, b.n, o)};
        b = better(b, x);
      }
      if (lane == 0) s_best[warp] = b;
      __syncthreads();
      if (warp == 0) {
        b = s_best[lane];
        for (int o = 16; o; o >>= 1) {
          BestRow x{__shfl_xor_sync(0xffffffffu, b.v, o), __shfl_xor_sync(0xffffffffu, b.n, o)};
          b = better(b, x);
        }
        if (lane == 0) {
          s_seed_n[ns] = b.n;
          s_seed_c[ns] = c;
          s_seed_p[ns] = b.n < 0 ? -FLT_MAX : b.v;
          s_num_seeds = ns + 1;
        }
      }
      __syncthreads();
    }
  }
  const int ns = s_num_seeds;
  // every RoI: the seed of highest IoU, first maximum (:125-133); eligibility bits for the capped pass
  for (int n = tid; n < R; n += blockDim.x) {
    float best = -FLT_MAX;
    int bi = -1;
    for (int i = 0; i < ns; i++) {
      const int g = s_seed_n[i];
      if (g < 0) continue;
      const float u = U[(size_t)n * R + g];
      if (best < u) {
        best = u;
        bi = i;
      }
    }
    unsigned char e = 0;
    int c = 0;
    if (bi >= 0) {
      c = s_seed_c[bi];
      e = (best >= fg ? 1 : 0) | ((best >= bg_lo && best < bg_hi) ? 2 : 0);
      RW[n] = CW ? CW[c] : s_seed_p[bi];  // provisional: the capped pass zeroes the rows that are not accepted
    } else {
      RW[n] = 0.f;
    }
    elig[n] = e;
    cls_tmp[n] = c;
  }
  if (tid < 2) s_carry[tid] = 0;
  __syncthreads();
  // caps in visiting order (:122-165): foreground while fewer than num_pos + 1 were accepted, else background
  // (if eligible) while fewer than num_neg + 1 were accepted, else weight 0.  Accepted-before counts = exclusive prefix
  // counts of the eligible ones, because every eligible row is accepted until the cap is hit.
  float wfg = 0.f, wbg = 0.f;
  for (int base = 0; base < R; base += blockDim.x) {
    const int p = base + tid;
    const int n = p < R ? (perm ? perm[p] : p) : -1;
    const unsigned char e = n >= 0 ? elig[n] : 0;
    const int f = e & 1;
    int incl = f;
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_scan[0][warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int v = s_scan[0][lane];
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
      }
      s_scan[0][lane] = v;
    }
    __syncthreads();
    const int f_before = s_carry[0] + (warp ? s_scan[0][warp - 1] : 0) + incl - f;
    const bool acc_fg = f && f_before <= num_pos;
    const int q = (!acc_fg && (e & 2)) ? 1 : 0;
    int incq = q;
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incq, o);
      if (lane >= o) incq += v;
    }
    if (lane == 31) s_scan[1][warp] = incq;
    __syncthreads();
    if (warp == 0) {
      int v = s_scan[1][lane];
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
      }
      s_scan[1][lane] = v;
    }
    __syncthreads();
    const int q_before = s_carry[1] + (warp ? s_scan[1][warp - 1] : 0) + incq - q;
    const bool acc_bg = q && q_before <= num_neg;
    if (n >= 0) {
      const float w = RW[n];
      RL[n] = acc_bg ? K : cls_tmp[n];
      if (acc_fg) wfg += w;
      else if (acc_bg) wbg += w;
      else RW[n] = 0.f;
    }
    __syncthreads();
    if (tid == 0) {
      s_carry[0] += s_scan[0][31];
      s_carry[1] += s_scan[1][31];
    }
    __syncthreads();
  }
  // call statistics (display only in the reference; fixed-order sums here)
  for (int o = 16; o; o >>= 1) {
    wfg += __shfl_xor_sync(0xffffffffu, wfg, o);
    wbg += __shfl_xor_sync(0xffffffffu, wbg, o);
  }
  if (lane == 0) {
    s_wsum[0][warp] = wfg;
    s_wsum[1][warp] = wbg;
  }
  __syncthreads();
  if (tid == 0) {
    float a = 0.f, b = 0.f;
    for (int i = 0; i < 32; i++) {
      a += s_wsum[0][i];
      b += s_wsum[1][i];
    }
    stats[0] = (float)min(s_carry[0], num_pos + 1);
    stats[1] = (float)min(s_carry[1], num_neg + 1);
    stats[2] = a;
    stats[3] = b;
  }
}

}  // namespace
}  // namespace jtsm

using namespace jtsm;

extern "C" size_t jtsm_roi_merge_workspace_bytes(int num_rois) {
  if (num_rois < 0) return 0;
  const int n = num_rois < 1 ? 1 : num_rois;
  return 256 + 4 * align_up((size_t)n * 4, 256) + align_up((size_t)kMergeMaxTop * kMergeMaxWindow * 4, 256) +
         align_up(merge_cub_bytes(n), 256);
}

extern "C" int jtsm_roi_merge_forward(const float* scores, const float* iou, const float* C, const float* D,
                                      int num_rois, int num_classes, float lambda, int top_k, int window, float* MC,
                                      float* MD, int* I, int* IC, int* stats, void* workspace, size_t workspace_bytes,
                                      void* stream) {
  if (num_rois < 0 || num_classes <= 0 || top_k < 1 || window < 1) return JTSM_ERR_INVALID_ARGUMENT;
  if (top_k > kMergeMaxTop || window > kMergeMaxWindow) return JTSM_ERR_UNSUPPORTED;
  if (!stats) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  if (num_rois == 0) {
    cudaMemsetAsync(stats, 0, 4 * sizeof(int), st);
    return cuda_status();
  }
  if (!scores || !iou || !C || !D || !MC || !MD || !I || !IC || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_roi_merge_workspace_bytes(num_rois)) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  MergeWs ws = carve_merge_ws(workspace, num_rois);
  iota_kernel<<<(num_rois + 255) / 256, 256, 0, st>>>(ws.vals_in, num_rois);
  cub::DeviceRadixSort::SortPairsDescending(ws.cub_tmp, ws.cub_bytes, scores, ws.keys_out, ws.vals_in, ws.order, num_rois,
                                            0, 32, st);
  cudaMemsetAsync(I, 0xFF, (size_t)num_rois * sizeof(int), st);
  roi_merge_clique_kernel<<<1, 1024, 0, st>>>(iou, ws.order, num_rois, lambda, top_k, window, I, IC, ws.members, ws.owner,
                                              stats);
  const long long total = (long long)num_rois * num_classes;
  int grid = (int)((total + 255) / 256);
  grid = grid > 4 * num_sms() ? 4 * num_sms() : grid;
  roi_merge_sum_kernel<<<grid, 256, 0, st>>>(C, D, num_rois, num_classes, IC, ws.members, ws.owner, stats, MC, MD);
  count_launch(3);
  return cuda_status();
}

extern "C" int jtsm_roi_merge_backward(const float* GMC, const float* GMD, const int* I, const int* IC, int num_rois,
                                       int num_classes, float* GC, float* GD, void* stream) {
  if (num_rois < 0 || num_classes <= 0) return JTSM_ERR_INVALID_ARGUMENT;
  if (num_rois == 0) return JTSM_OK;
  if (!GMC || !GMD || !I || !IC || !GC || !GD) return JTSM_ERR_INVALID_ARGUMENT;
  const long long total = (long long)num_rois * num_classes;
  int grid = (int)((total + 255) / 256);
  grid = grid > 4 * num_sms() ? 4 * num_sms() : grid;
  roi_merge_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(GMC, GMD, I, IC, num_rois, num_classes, GC, GD);
  count_launch();
  return cuda_status();
}

extern "C" size_t jtsm_roi_label_workspace_bytes(int num_rois) {
  if (num_rois < 0) return 0;
  return 512 + align_up((size_t)num_rois, 256) + (size_t)num_rois * 4;
}

extern "C" int jtsm_roi_label_forward(const float* S, int score_stride, const float* U, const float* L, const float* CW,
                                      const int* perm, int num_rois, int num_classes, float fg_thresh,
                                      float bg_thresh_hi, float bg_thresh_lo, int num_pos, int num_neg, int top_k, int* RL,
                                      float* RW, float* stats, void* workspace, size_t workspace_bytes, void* stream) {
  if (num_rois < 0 || num_classes <= 0 || score_stride < num_classes || top_k < 1 || num_pos < 0 || num_neg < 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (!stats) return JTSM_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  if (num_rois == 0) {
    cudaMemsetAsync(stats, 0, 4 * sizeof(float), st);
    return cuda_status();
  }
  if (!S || !U || !L || !RL || !RW || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_roi_label_workspace_bytes(num_rois)) return JTSM_ERR_WORKSPACE_TOO_SMALL;
  char* p = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  unsigned char* elig = (unsigned char*)p;
  int* cls_tmp = (int*)(p + align_up((size_t)num_rois, 256));
  roi_label_kernel<<<1, 1024, 0, st>>>(S, score_stride, U, L, CW, perm, num_rois, num_classes, fg_thresh, bg_thresh_hi,
                                       bg_thresh_lo, num_pos, num_neg, top_k, RL, RW, stats, elig, cls_tmp);
  count_launch();
  return cuda_status();
}
