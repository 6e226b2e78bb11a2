// engine.h — per-device state of the probe engine (host side of libgpuprobe.so).
//
// One Device per CUDA-visible GPU: its own stream, event pool, scratch block,
// pinned result buffer and a grow-only arena that every probe carves its
// arrays from.  Reference call sites that feed this: device/device_map.go:48-75
// (one visit per GPU) and plugin/manager.go:156-174 (loadPlugins).
#pragma once
#include <cuda_runtime.h>

#include <cstdlib>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gpuprobe.h"
#include "kernels.h"

namespace gp {

constexpr int kEventPool = 320;
constexpr int kMaxResults = 256;

struct Device {
  int ordinal = -1;
  char uuid[48] = {0};
  int sm_count = 0;
  int cc_major = 0, cc_minor = 0;
  uint64_t total_mem = 0;

  std::mutex mu;
  bool ready = false;
  cudaStream_t stream = nullptr;
  uint8_t* arena = nullptr;
  size_t arena_bytes = 0;
  // VMM arena (arena.cu): one reserved virtual range, physical chunks mapped from its base upwards
  bool arena_is_vmm = false;
  unsigned long long va_base = 0;           // CUdeviceptr of the reservation, 0 until first use
  size_t va_bytes = 0;
  struct Chunk { unsigned long long handle; size_t bytes; };
  std::vector<Chunk> chunks;
  unsigned long long peer_mask = 0;         // ordinals (bit i) that have been granted access to the mapped range
  double alloc_s = 0, free_s = 0;           // host seconds spent growing / releasing since the caller last zeroed them
  std::thread reaper;                       // deferred release of the last arena (free_arena_deferred)
  bool poisoned = false;                    // a sticky CUDA fault was seen: context dead until gp_device_reset
  int poison_code = 0;                      // the cudaError_t that did it
  uint8_t* scratch = nullptr;               // kVerifyScratchBytes | results | counters
  gp_pattern_result* d_results = nullptr;   // [kMaxResults]
  unsigned long long* d_count = nullptr;    // triad-check counter
  gp_pattern_result* h_results = nullptr;   // pinned, [kMaxResults]
  unsigned long long* h_count = nullptr;    // pinned
  cudaEvent_t events[kEventPool] = {};
  gp_kernel_stat stats[GP_K_COUNT] = {};
  uint64_t triad_operands_bytes = 0;        // b, c generated for this size at arena+3N / +4N
  uint64_t triad_operands_seed = 0;
};

// host-side fp64 reference of the GEMM probe's sampled entries: a pure function of (seed, shape, sample count),
// so it is computed once per context and shared by every device and every later probe (gemm.cu)
struct GemmRef {
  uint64_t seed = 0;
  uint32_t m = 0, n = 0, k = 0, samples = 0;
  std::vector<uint32_t> ij;        // 2 * samples: row, column
  std::vector<double> acc, aacc;   // sum_k a*b and sum_k |a*b| per sample
};

struct Ctx {
  gp_config cfg;
  std::mutex gemm_ref_mu;
  std::vector<std::shared_ptr<GemmRef>> gemm_refs;
  std::vector<std::unique_ptr<Device>> devs;
  std::mutex multi_mu;  // serialises the cross-device probes (P2P, NCCL)
  void* nccl_handle = nullptr;
};

// Development switches (launch shapes, raster, cache hints) are environment variables only in builds made
// with -DGPUPROBE_DEV_SWITCHES (tools/); a release library reads none of them, so a stray variable in a
// DaemonSet's environment cannot change what the probe measures.
#ifdef GPUPROBE_DEV_SWITCHES
#define GP_DEV_SWITCH(name, dflt) (getenv(name) ? atoi(getenv(name)) : (dflt))
#else
#define GP_DEV_SWITCH(name, dflt) (dflt)
#endif

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, int64_t* detail);
// true for the CUDA errors that kill the context (every later call returns them again)
bool cuda_error_is_sticky(cudaError_t e);
// ordinals (bit i) on which cuda_fail has seen a sticky error since the last clear
unsigned long long poisoned_ordinals();
void clear_poisoned_ordinal(int ordinal);

Device* find_device(Ctx* ctx, const char* uuid);
int device_prepare(Ctx* ctx, Device* d);                  // lazy per-device init, lock held
int ensure_arena(Ctx* ctx, Device* d, size_t bytes, bool ignore_cap = false);  // lock held
void free_arena(Device* d);
void free_arena_deferred(Device* d);         // same, on a reaper thread; returns at once
void free_buffers_deferred(Device* d, const std::vector<void*>& ptrs);
void arena_join_reaper(Device* d);
void release_arena_va(Device* d);            // at shutdown / reset: give the virtual range back too
// make d's mapped arena readable and writable from the devices in `ordinal_mask` (VMM memory is not
// covered by cudaDeviceEnablePeerAccess); a no-op for a cudaMalloc arena
int arena_grant_peers(Device* d, unsigned long long ordinal_mask);
uint64_t cfg_arena_cap(const Ctx* c);
// noalloc.cu: forget the cached symbol address of a device (its context is being reset)
void noalloc_forget_device(int ordinal);
int resolve_variant(const Ctx* ctx);

// timing of `iters` launches with one event between consecutive launches
struct Timed {
  std::vector<float> ms;
};
void summarize(const std::vector<float>& ms, gp_timing* t);

// implemented in gemm.cu / p2p.cu
int gemm_probe_locked(Ctx* ctx, Device* d, uint32_t mnk, uint64_t seed, uint32_t warmup,
                      uint32_t iters, uint32_t samples, gp_gemm_result* out, uint8_t* workspace = nullptr,
                      const std::function<void()>* after_timed = nullptr);
size_t gemm_probe_bytes(const Ctx* ctx, uint32_t mnk, uint32_t samples);
int gemm_host_locked(Ctx* ctx, Device* d, const uint16_t* a, const uint16_t* b, float* c,
                     uint32_t m, uint32_t n, uint32_t k);
int p2p_probe(Ctx* ctx, int ngpus, size_t bytes, int dir, double* gbps);
int fabric_probe(Ctx* ctx, int ngpus, gp_fabric_result* out);
int allreduce_probe(Ctx* ctx, int nranks, size_t bytes, uint32_t warmup, uint32_t iters,
                    double* algbw, double* busbw);
void nccl_teardown(Ctx* ctx);

}  // namespace gp

struct gp_ctx : gp::Ctx {};
