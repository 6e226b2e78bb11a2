// engine.h — per-device state of the probe engine (host side of libgpuprobe.so).
//
// One Device per CUDA-visible GPU: its own stream, event pool, scratch block,
// pinned result buffer and a grow-only arena that every probe carves its
// arrays from.  Reference call sites that feed this: device/device_map.go:48-75
// (one visit per GPU) and plugin/manager.go:156-174 (loadPlugins).
#pragma once
#include <cuda_runtime.h>

#include <memory>
#include <mutex>
#include <string>
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

struct Ctx {
  gp_config cfg;
  std::vector<std::unique_ptr<Device>> devs;
  std::mutex multi_mu;  // serialises the cross-device probes (P2P, NCCL)
  void* nccl_handle = nullptr;
};

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, int64_t* detail);

Device* find_device(Ctx* ctx, const char* uuid);
int device_prepare(Ctx* ctx, Device* d);                  // lazy per-device init, lock held
int ensure_arena(Ctx* ctx, Device* d, size_t bytes, bool ignore_cap = false);  // lock held
void free_arena(Device* d);
int resolve_variant(const Ctx* ctx);

// timing of `iters` launches with one event between consecutive launches
struct Timed {
  std::vector<float> ms;
};
void summarize(const std::vector<float>& ms, gp_timing* t);

// implemented in gemm.cu / p2p.cu
int gemm_probe_locked(Ctx* ctx, Device* d, uint32_t mnk, uint64_t seed, uint32_t warmup,
                      uint32_t iters, uint32_t samples, gp_gemm_result* out);
int gemm_host_locked(Ctx* ctx, Device* d, const uint16_t* a, const uint16_t* b, float* c,
                     uint32_t m, uint32_t n, uint32_t k);
int p2p_probe(Ctx* ctx, int ngpus, size_t bytes, int dir, double* gbps);
int allreduce_probe(Ctx* ctx, int nranks, size_t bytes, uint32_t warmup, uint32_t iters,
                    double* algbw, double* busbw);
void nccl_teardown(Ctx* ctx);

}  // namespace gp

struct gp_ctx : gp::Ctx {};
