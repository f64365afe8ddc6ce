// RoIPool -> fc6 fused on the 5th-generation tensor cores (inference first: no argmax is kept).
//
// Reference: the pooled [R, C, P, P] tensor is materialised, scaled by (objectness + 1), flattened and fed to fc6
// (projects/WSL/wsl/modeling/roi_heads/roi_heads_oicr.py:391-403, box_head.py:59-68,90):
//     out = relu(fc6(flatten(roi_pool(x, rois) * (objectness + 1))))            200 MB written + read at config 2
// Here the pooled values only ever exist as A-operand tiles in shared memory:
//     out[r, n] = relu(bias[n] + sum_k A[r, k] * Wp[n, k]),  k = bin * C + c  (bin-major: 64 consecutive k are 64
//     consecutive CHANNELS of one bin, i.e. one 128-byte run of an NHWC bf16 feature map per cell; fc6's weight is
//     re-laid out once: Wp[n, bin * C + c] = W[n, c * P*P + bin])
// One CTA owns a 128-RoI x 512-column output tile (the whole TMEM: 128 lanes x 512 fp32 columns) and walks the
// K = P*P*C dimension in slices of 64:
//   warps 0-3 (128 threads, thread = RoI row): scan the RoI's bin with 16-byte channel-contiguous loads (__hmax2),
//              scale, write the row's 128 bytes into the A stage in the canonical 128B-swizzled K-major layout,
//              fence.proxy.async, arrive on full_a[stage]
//   warp 4 lane 0: TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) of the 512 x 64 weight tile into the B stage
//   warp 5 lane 0: tcgen05.mma.cta_group::1.kind::f16 (M 128, N 256, K 16) x 4 x 2 per slice, accumulators in TMEM;
//              tcgen05.commit frees the stage, the last commit signals the epilogue
//   warps 0-3: tcgen05.ld 32x32b -> + bias, ReLU -> global (fp32)
// bf16 operands, fp32 accumulation.  This is the PROTOTYPE the round-1 review asked for: every CTA along N re-gathers
// the same A tile (8x redundant pooling at N = 4096); DESIGN.md section 8 has the measured comparison with
// (RoIPool kernel + cuBLAS bf16 GEMM) and what a cluster-multicast A tile would change.
#include <cuda.h>  // CUtensorMap (types only; cuTensorMapEncodeTiled is resolved at run time, no libcuda link)

#include "roi_bins.cuh"
#include "roi_pool_rmqk.cuh"  // mbarrier helpers

namespace jtsm {

constexpr int kFcM = 128;         // RoIs per CTA (TMEM lanes)
constexpr int kFcN = 512;         // output columns per CTA (TMEM columns)
constexpr int kFcK = 64;          // K per slice: 64 bf16 = one 128-byte swizzle row
constexpr int kFcStages = 2;
constexpr int kFcThreads = 192;   // 4 producer / epilogue warps + TMA warp + MMA warp
constexpr int kFcABytes = kFcM * kFcK * 2;   // 16 KB
constexpr int kFcBBytes = kFcN * kFcK * 2;   // 64 KB (two 256-row TMA boxes)
constexpr size_t kFcSmem = 1024 /* alignment slack */ + (size_t)kFcStages * (kFcABytes + kFcBBytes) + 256;

__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int x, int y, unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y)
      : "memory");
}
// shared-memory matrix descriptor, K-major, 128-byte swizzle: 8-row groups 1024 B apart (SBO), version 1 (sm_100)
__device__ __forceinline__ unsigned long long umma_desc_sw128(const void* smem) {
  unsigned long long d = 0;
  d |= (unsigned long long)((smem_u32(smem) & 0x3FFFFu) >> 4);  // start address
  d |= (unsigned long long)1 << 16;                              // leading byte offset (unused for swizzled K-major)
  d |= (unsigned long long)(1024 >> 4) << 32;                    // stride byte offset
  d |= (unsigned long long)1 << 46;                              // descriptor version
  d |= (unsigned long long)2 << 61;                              // SWIZZLE_128B
  return d;
}
// instruction descriptor kind::f16: D fp32, A / B bf16, both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
__device__ __forceinline__ unsigned umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(unsigned d_tmem, unsigned long long a_desc, unsigned long long b_desc,
                                          unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

__global__ void __launch_bounds__(kFcThreads, 1)
pool_fc6_kernel(const __nv_bfloat16* __restrict__ feat /* NHWC */, const __grid_constant__ CUtensorMap wmap,
                const float* __restrict__ bias, int channels, int H, int W, int PH, int PW, int num_rois,
                int out_features, int relu, PoolWorkspace ws, float* __restrict__ out) {
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* base = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);  // SW128 tiles: 1024 B aligned
  unsigned char* a_tiles = base;
  unsigned char* b_tiles = base + kFcStages * kFcABytes;
  unsigned long long* bars = (unsigned long long*)(b_tiles + kFcStages * kFcBBytes);
  unsigned long long* full_a = bars;                 // [stages] 128 producer arrivals
  unsigned long long* full_b = bars + kFcStages;     // [stages] 1 arrival + TMA bytes
  unsigned long long* empty = bars + 2 * kFcStages;  // [stages] 1 arrival (tcgen05.commit)
  unsigned long long* accum = bars + 3 * kFcStages;  // accumulator complete
  unsigned* tmem_slot = (unsigned*)(bars + 3 * kFcStages + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kFcM, n0 = blockIdx.y * kFcN;
  const int cgroups = channels / kFcK, PP = PH * PW;
  const int slices = PP * cgroups;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kFcStages; ++s) {
      mbar_init(full_a + s, kFcM);
      mbar_init(full_b + s, 1);
      mbar_init(empty + s, 1);
    }
    mbar_init(accum, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) {  // TMEM: all 512 columns; the allocating warp also frees them
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned tmem = *tmem_slot;

  if (warp < 4) {
    // ===== A producer: thread = RoI row
    const int row = threadIdx.x, r = m0 + row;
    const bool live = r < num_rois && ws.batch_idx[min(r, num_rois - 1)] >= 0;
    const int n = live ? ws.batch_idx[r] : 0;
    const float scale = live ? ws.rscale[r] : 0.f;
    const __nv_bfloat16* fmap = feat + (size_t)n * H * W * channels;
    for (int s = 0; s < slices; ++s) {
      const int st = s % kFcStages, it = s / kFcStages;
      const int bin = s / cgroups, c0 = (s - bin * cgroups) * kFcK;
      const int ph = bin / PW, pw = bin - ph * PW;
      __nv_bfloat162 acc[32];
      const __nv_bfloat162 ninf = __float2bfloat162_rn(-INFINITY);
#pragma unroll
      for (int j = 0; j < 32; ++j) acc[j] = ninf;
      bool any = false;
      if (live) {
        const uint32_t hb = __ldg(ws.hbins + (size_t)r * PH + ph), wb = __ldg(ws.wbins + (size_t)r * PW + pw);
        const int hs = range_lo(hb), he = range_hi(hb), wst = range_lo(wb), wen = range_hi(wb);
        any = he > hs && wen > wst;
        for (int h = hs; h < he; ++h)
          for (int w = wst; w < wen; ++w) {
            const uint4* cell = (const uint4*)(fmap + ((size_t)h * W + w) * channels + c0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint4 v = __ldg(cell + j);
              acc[4 * j + 0] = __hmax2(acc[4 * j + 0], *(const __nv_bfloat162*)&v.x);
              acc[4 * j + 1] = __hmax2(acc[4 * j + 1], *(const __nv_bfloat162*)&v.y);
              acc[4 * j + 2] = __hmax2(acc[4 * j + 2], *(const __nv_bfloat162*)&v.z);
              acc[4 * j + 3] = __hmax2(acc[4 * j + 3], *(const __nv_bfloat162*)&v.w);
            }
          }
      }
      if (it > 0) mbar_wait(empty + st, (unsigned)((it - 1) & 1));  // the MMAs that read this stage have completed
      unsigned char* arow = a_tiles + st * kFcABytes + row * 128;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint4 o;
        unsigned* ow = &o.x;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float2 f = any ? __bfloat1622float2(acc[4 * j + q]) : make_float2(0.f, 0.f);
          const __nv_bfloat162 b = __floats2bfloat162_rn(__fmul_rn(f.x, scale), __fmul_rn(f.y, scale));
          ow[q] = *(const unsigned*)&b;
        }
        *(uint4*)(arow + ((j ^ (row & 7)) << 4)) = o;  // 128-byte swizzle: 16-byte chunk index XOR (row mod 8)
      }
      fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's async proxy
      mbar_arrive(full_a + st);
    }
    // ===== epilogue: TMEM -> registers -> + bias, ReLU -> global
    mbar_wait(accum, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned lane_base = tmem + ((unsigned)(warp * 32) << 16);
    for (int cb = 0; cb < kFcN; cb += 32) {
      unsigned v[32];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
            "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
            "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(lane_base + (unsigned)cb));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (r < num_rois) {
        float* orow = out + (size_t)r * out_features + n0 + cb;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          float4 o;
          o.x = __uint_as_float(v[j + 0]) + __ldg(bias + n0 + cb + j + 0);
          o.y = __uint_as_float(v[j + 1]) + __ldg(bias + n0 + cb + j + 1);
          o.z = __uint_as_float(v[j + 2]) + __ldg(bias + n0 + cb + j + 2);
          o.w = __uint_as_float(v[j + 3]) + __ldg(bias + n0 + cb + j + 3);
          if (relu) o = make_float4(fmaxf(o.x, 0.f), fmaxf(o.y, 0.f), fmaxf(o.z, 0.f), fmaxf(o.w, 0.f));
          *(float4*)(orow + j) = o;
        }
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  } else if (warp == 4) {
    // ===== TMA producer of the weight tiles (one lane)
    if (lane == 0) {
      for (int s = 0; s < slices; ++s) {
        const int st = s % kFcStages, it = s / kFcStages;
        if (it > 0) mbar_wait(empty + st, (unsigned)((it - 1) & 1));
        mbar_expect_tx(full_b + st, kFcBBytes);
        unsigned char* bt = b_tiles + st * kFcBBytes;
        tma_load_2d(bt, &wmap, s * kFcK, n0, full_b + st);
        tma_load_2d(bt + kFcBBytes / 2, &wmap, s * kFcK, n0 + 256, full_b + st);
      }
    }
  } else {
    // ===== MMA issuer (one lane)
    if (lane == 0) {
      const unsigned idesc = umma_idesc_bf16(kFcM, 256);
      for (int s = 0; s < slices; ++s) {
        const int st = s % kFcStages, it = s / kFcStages;
        mbar_wait(full_a + st, (unsigned)(it & 1));
        mbar_wait(full_b + st, (unsigned)(it & 1));
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const unsigned long long ad = umma_desc_sw128(a_tiles + st * kFcABytes);
        const unsigned long long bd0 = umma_desc_sw128(b_tiles + st * kFcBBytes);
        const unsigned long long bd1 = umma_desc_sw128(b_tiles + st * kFcBBytes + kFcBBytes / 2);
#pragma unroll
        for (int k = 0; k < kFcK / 16; ++k) {  // K = 16 per instruction: 32 bytes along the swizzled row -> +2 in the address field
          const unsigned acc = (s > 0 || k > 0) ? 1u : 0u;
          umma_bf16(tmem, ad + 2ull * k, bd0 + 2ull * k, idesc, acc);
          umma_bf16(tmem + 256u, ad + 2ull * k, bd1 + 2ull * k, idesc, acc);
        }
        umma_commit(empty + st);  // arrives when the MMAs above have read the stage (implies fence::before_thread_sync)
      }
      umma_commit(accum);
    }
  }
  __syncthreads();
  if (warp == 4) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

}  // namespace jtsm

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
using namespace jtsm;

namespace jtsm {
int moi_rmq_prep(const float* rois, int num_rois, int batch, int channels, int height, int width, float scale, int PH,
                 int PW, const float* row_scale, const PoolWorkspace& ws, cudaStream_t st);
}

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
  static std::atomic<void*> cached{nullptr};
  void* fn = cached.load(std::memory_order_acquire);
  if (!fn) {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    cached.store(fn, std::memory_order_release);
  }
  return (EncodeTiledFn)fn;
}
}  // namespace

extern "C" size_t jtsm_pool_fc6_workspace_bytes(int batch, int num_rois, int pooled_h, int pooled_w) {
  if (batch < 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0) return 0;
  return pool_workspace_bytes(batch, num_rois, pooled_h, pooled_w, 1);
}

extern "C" int jtsm_pool_fc6_forward(const void* feat_nhwc_bf16, const float* rois, int batch, int channels, int height,
                                     int width, int num_rois, float spatial_scale, int pooled_h, int pooled_w,
                                     const float* row_scale, const void* weight_binmajor_bf16, const float* bias,
                                     int out_features, int relu, float* out, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  if (batch <= 0 || channels <= 0 || height <= 0 || width <= 0 || num_rois < 0 || pooled_h <= 0 || pooled_w <= 0 ||
      out_features <= 0)
    return JTSM_ERR_INVALID_ARGUMENT;
  if (channels % kFcK != 0 || out_features % kFcN != 0) return JTSM_ERR_UNSUPPORTED;  // prototype tile sizes
  if (height > 32767 || width > 32767) return JTSM_ERR_UNSUPPORTED;
  if (num_rois == 0) return JTSM_OK;
  if (!feat_nhwc_bf16 || !rois || !weight_binmajor_bf16 || !bias || !out || !workspace) return JTSM_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)feat_nhwc_bf16 | (uintptr_t)weight_binmajor_bf16 | (uintptr_t)out) & 15) return JTSM_ERR_INVALID_ARGUMENT;
  if (workspace_bytes < jtsm_pool_fc6_workspace_bytes(batch, num_rois, pooled_h, pooled_w))
    return JTSM_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  PoolWorkspace ws = carve_pool_workspace(workspace, batch, num_rois, pooled_h, pooled_w, 1);
  int rc = moi_rmq_prep(rois, num_rois, batch, channels, height, width, spatial_scale, pooled_h, pooled_w, row_scale, ws, st);
  if (rc != JTSM_OK) return rc;
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return JTSM_ERR_UNSUPPORTED;
  // weights [out_features, K] bf16 row-major, K = P*P*C; box = 64 (K) x 256 (rows), 128-byte swizzle
  const long long K = (long long)pooled_h * pooled_w * channels;
  CUtensorMap wmap;
  const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)out_features};
  const cuuint64_t gstride[1] = {(cuuint64_t)K * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kFcK, 256u};
  const cuuint32_t estr[2] = {1u, 1u};
  if (enc(&wmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(weight_binmajor_bf16), gdim, gstride, box, estr,
          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return JTSM_ERR_INVALID_ARGUMENT;
  cudaFuncSetAttribute(pool_fc6_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFcSmem);
  dim3 grid((num_rois + kFcM - 1) / kFcM, out_features / kFcN);
  pool_fc6_kernel<<<grid, kFcThreads, kFcSmem, st>>>((const __nv_bfloat16*)feat_nhwc_bf16, wmap, bias, channels, height,
                                                     width, pooled_h, pooled_w, num_rois, out_features, relu, ws, out);
  count_launch();
  return cuda_status();
}
