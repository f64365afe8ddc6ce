// Library-wide state: ABI version, error strings, launch counter, tuning knobs.
#include <string.h>

#include <map>
#include <mutex>
#include <string>

#include "common.cuh"

namespace jtsm {

std::atomic<unsigned long long> g_launch_count{0};

namespace {
std::mutex g_tuning_mu;
std::map<std::string, int>& tuning_map() {
  static std::map<std::string, int> m = {
      {"roi_pool_splits", 0},      // 0 = auto
      {"roi_pool_force_gmem", 0},  // 1 = skip the smem-plane kernel
      {"roi_pool_bwd_algo", 0},    // 0 = auto
      {"roi_pool_fwd_algo", 0},    // 0 = auto, 1 = direct smem scan, 2 = gmem scan, 3 = range-max tables
      {"roi_pool_fwd_async", 1},   // 1 = fetch the next plane with cp.async when a second plane buffer fits
      {"roi_align_bwd_algo", 0},   // 0 = auto (separable: column kernel for PH <= 16, else cell kernel), 1 = one RED per tap, 2 = cell kernel
      {"roi_align_fwd_algo", 0},   // 0 = auto (tap tables), 1 = coordinates recomputed per sample
      {"roi_align_force_gmem", 0},
      {"roi_align_fwd_pairs", 1},  // 1 = staged plane as (x, x+1) pairs: one 8-byte shared load per row of a sample
      {"moi_rmq", 60},             // MOIPool: fully switched-on bins go to the RoIPool range-max kernel when <= this % of the bins are partially masked (device-side gate); -1 = off
      {"loop_algo", 0},            // ROILoopPool forward: 0 = auto (virtual RoIs: range-max + visit lists), 1 = direct scans
      {"moi_planes", 2},           // channels per CTA of the MOIPool visit-list kernel (1 or 2)
      {"moi_algo", 0},             // MOIPool forward: 0 = auto (per-bin visit lists), 1 = row bit masks, 2 = byte-mask scan
  };
  return m;
}
}  // namespace

int tuning(const char* key, int dflt) {
  std::lock_guard<std::mutex> lk(g_tuning_mu);
  auto& m = tuning_map();
  auto it = m.find(key);
  return it == m.end() ? dflt : it->second;
}

int num_sms() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      cached = kNumSMs;
  }
  return cached;
}

}  // namespace jtsm

extern "C" int jtsm_abi_version(void) { return JTSM_ABI_VERSION; }

extern "C" const char* jtsm_error_string(int code) {
  switch (code) {
    case JTSM_OK: return "ok";
    case JTSM_ERR_INVALID_ARGUMENT: return "invalid argument";
    case JTSM_ERR_UNSUPPORTED: return "unsupported configuration";
    case JTSM_ERR_WORKSPACE_TOO_SMALL: return "workspace too small";
    default: break;
  }
  if (code >= JTSM_ERR_CUDA_BASE) return cudaGetErrorString((cudaError_t)(code - JTSM_ERR_CUDA_BASE));
  return "unknown error";
}

extern "C" unsigned long long jtsm_launch_count(void) {
  return jtsm::g_launch_count.load(std::memory_order_relaxed);
}

extern "C" int jtsm_set_tuning(const char* key, int value) {
  if (!key) return JTSM_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(jtsm::g_tuning_mu);
  auto& m = jtsm::tuning_map();
  auto it = m.find(key);
  if (it == m.end()) return JTSM_ERR_INVALID_ARGUMENT;
  it->second = value;
  return JTSM_OK;
}

extern "C" int jtsm_get_tuning(const char* key, int* value) {
  if (!key || !value) return JTSM_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(jtsm::g_tuning_mu);
  auto& m = jtsm::tuning_map();
  auto it = m.find(key);
  if (it == m.end()) return JTSM_ERR_INVALID_ARGUMENT;
  *value = it->second;
  return JTSM_OK;
}
