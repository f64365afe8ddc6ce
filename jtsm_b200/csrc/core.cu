// Library-wide state: ABI version, error strings, launch counter, tuning knobs.
#include <string.h>


#include "common.cuh"

namespace jtsm {

std::atomic<unsigned long long> g_launch_count{0};

namespace {
// Tuning knobs: a fixed table of atomics (no lock, no allocation on the launch path); a launcher resolves a key by
// comparing at most kNumKnobs short strings, and the hot launchers cache the slot index in a function-local static.
struct Knob {
  const char* key;
  std::atomic<int> value;
};
Knob g_knobs[] = {
    {"roi_pool_splits", {0}},      // 0 = auto
    {"roi_pool_force_gmem", {0}},  // 1 = skip the smem-plane kernel
    {"roi_pool_fwd_rows", {0}},       // range-max kernel: row-mapped main loop (a thread keeps one bin position); opt-in
    {"roi_pool_fwd_bandlist", {1}},   // banded planes: per-band RoI lists (1) or the band filter over all bins (0)
    {"roi_pool_fwd_pack16", {1}},     // bf16 / fp16 maps: packed (value, index) tables, 4 look-ups per bin (roi_pool_rmqp.cuh); 1 = single-band planes, 2 = banded too
    {"roi_pool_bwd_super_mb", {0}},   // 7 x 7 backward: MB of accumulation planes the warps in flight spread over (0 = auto)
    {"roi_pool_bwd_algo", {0}},    // 0 = auto (7 x 7: all 49 bins of a (RoI, channel) per lane, else vec4 folding), 1 = one RED per element, 3 = vec4 folding
    {"roi_pool_fwd_algo", {0}},    // 0 = auto, 1 = direct smem scan, 2 = gmem scan, 3 = range-max tables
    {"roi_pool_fwd_keys", {0}},    // range-max tables: 0 = absolute indices, 8 look-ups per bin (default); 1 = (key10 | rel6) entries, 5 look-ups, ALU bound
    {"roi_pool_fwd_tma", {1}},     // 1 = planes arrive by cp.async.bulk (TMA 1-D) + mbarrier into a staging buffer; 0 = plain loads
    {"roi_align_bwd_algo", {0}},   // 0 = auto (separable: column kernel for PH <= 16, else cell kernel), 1 = one RED per tap, 2 = cell kernel
    {"roi_align_fwd_algo", {0}},   // 0 = auto (tap tables), 1 = coordinates recomputed per sample
    {"roi_align_force_gmem", {0}},
    {"roi_align_fwd_pairs", {1}},  // 1 = staged plane as (x, x+1) pairs: one 8-byte shared load per row of a sample
    {"moi_rmq", {60}},             // MOIPool: fully switched-on bins go to the RoIPool range-max kernel when <= this % of the bins are partially masked (device-side gate); -1 = off
    {"loop_algo", {0}},            // ROILoopPool forward: 0 = auto (virtual RoIs: range-max + visit lists), 1 = direct scans
    {"moi_planes", {2}},           // channels per CTA of the MOIPool visit-list kernel (1 or 2)
    {"moi_algo", {0}},             // MOIPool forward: 0 = auto (per-bin visit lists), 1 = row bit masks, 2 = byte-mask scan
    {"nms_algo", {0}},             // 0 = auto (cluster of CTAs per class when classes are few), 1 = one CTA per class
    {"infer_filter_algo", {0}},    // reserved
};
constexpr int kNumKnobs = (int)(sizeof(g_knobs) / sizeof(g_knobs[0]));
Knob* find_knob(const char* key) {
  for (int i = 0; i < kNumKnobs; ++i)
    if (strcmp(g_knobs[i].key, key) == 0) return &g_knobs[i];
  return nullptr;
}
}  // namespace

std::atomic<int>* find_knob_public(const char* key) {
  Knob* k = find_knob(key);
  return k ? &k->value : nullptr;
}

int tuning(const char* key, int dflt) {
  const Knob* k = find_knob(key);
  return k ? k->value.load(std::memory_order_relaxed) : dflt;
}

// SM count of the CURRENT device (cached per device ordinal: one process may drive several GPUs).
int num_sms() {
  constexpr int kMaxDev = 64;
  static std::atomic<int> cached[kMaxDev];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) return kNumSMs;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = kNumSMs;
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

}  // namespace jtsm

extern "C" int jtsm_abi_version(void) { return JTSM_ABI_VERSION; }

extern "C" const char* jtsm_error_string(int code) {
  switch (code) {
    case JTSM_OK: return "ok";
    case JTSM_ERR_INVALID_ARGUMENT: return "invalid argument";
    case JTSM_ERR_UNSUPPORTED: return "unsupported configuration";
    case JTSM_ERR_WORKSPACE_TOO_SMALL: return "workspace too small";
    case JTSM_ERR_STATIC_LIMIT: return "size exceeds the static scratch limit of this allocation-free entry point";
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
  auto* k = jtsm::find_knob_public(key);
  if (!k) return JTSM_ERR_INVALID_ARGUMENT;
  k->store(value, std::memory_order_relaxed);
  return JTSM_OK;
}

extern "C" int jtsm_get_tuning(const char* key, int* value) {
  if (!key || !value) return JTSM_ERR_INVALID_ARGUMENT;
  auto* k = jtsm::find_knob_public(key);
  if (!k) return JTSM_ERR_INVALID_ARGUMENT;
  *value = k->load(std::memory_order_relaxed);
  return JTSM_OK;
}
