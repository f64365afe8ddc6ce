// Shared helpers for the jtsm_b200 CUDA kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include <atomic>

#include "../../include/jtsm_b200.h"

namespace jtsm {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// --- host-side bookkeeping -------------------------------------------------
extern std::atomic<unsigned long long> g_launch_count;
inline void count_launch(int n = 1) { g_launch_count.fetch_add(n, std::memory_order_relaxed); }

int tuning(const char* key, int dflt);
int num_sms();  // SM count of the current device (cached)

inline int cuda_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? JTSM_OK : JTSM_ERR_CUDA_BASE + (int)e;
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// --- element type traits ---------------------------------------------------
template <typename T> struct Elem;
template <> struct Elem<float> {
  static __device__ __forceinline__ float to_f(float v) { return v; }
  static __device__ __forceinline__ float from_f(float v) { return v; }
};
template <> struct Elem<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
};
template <> struct Elem<__half> {
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
};

// argmax storage
template <typename A> struct Arg;
template <> struct Arg<int> {
  static __device__ __forceinline__ int enc(int i) { return i; }
  static __device__ __forceinline__ int dec(int v) { return v; }
};
template <> struct Arg<unsigned short> {
  static __device__ __forceinline__ unsigned short enc(int i) { return (unsigned short)(i < 0 ? 0xFFFF : i); }
  static __device__ __forceinline__ int dec(unsigned short v) { return v == 0xFFFF ? -1 : (int)v; }
};

// Dispatch helpers: call f(T{}) for the element type named by `dtype`.
#define JTSM_DISPATCH_DTYPE(dtype, T, ...)                       \
  switch (dtype) {                                               \
    case JTSM_F32: { using T = float; __VA_ARGS__; break; }      \
    case JTSM_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; } \
    case JTSM_F16: { using T = __half; __VA_ARGS__; break; }     \
    default: return JTSM_ERR_INVALID_ARGUMENT;                   \
  }
#define JTSM_DISPATCH_ARGMAX(adt, A, ...)                              \
  switch (adt) {                                                       \
    case JTSM_ARGMAX_I32: { using A = int; __VA_ARGS__; break; }       \
    case JTSM_ARGMAX_U16: { using A = unsigned short; __VA_ARGS__; break; } \
    default: return JTSM_ERR_INVALID_ARGUMENT;                         \
  }

// class-column register slots per lane (K <= 32*KT)
#define JTSM_DISPATCH_KT(K, KT, ...)                         \
  if ((K) <= 32) { constexpr int KT = 1; __VA_ARGS__; }      \
  else if ((K) <= 64) { constexpr int KT = 2; __VA_ARGS__; } \
  else if ((K) <= 128) { constexpr int KT = 4; __VA_ARGS__; }\
  else { constexpr int KT = 8; __VA_ARGS__; }

// --- warp helpers ----------------------------------------------------------
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace jtsm
