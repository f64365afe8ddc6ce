// Host definitions of the few CUDA names the reference's WSL pooling kernels use, so that the kernel bodies in
// projects/WSL/wsl/layers/csrc/{ROILoopPool,MOIPool}/*_cuda.cu compile with g++ and run serially on the CPU.
// TEST INFRASTRUCTURE ONLY (oracle/_ref): never linked into the product.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
using std::max;
using std::min;
template <typename T>
static inline T atomicAdd(T* addr, T v) {
  T old = *addr;
  *addr = old + v;
  return old;
}
