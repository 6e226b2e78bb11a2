// engine.cu — host side of libgpuprobe.so: device table, arena, timing, the
// per-device probe and the C ABI declared in include/gpuprobe.h.
//
// What the reference does at these points today: nothing.  BuildDevice
// (device/devices.go:41-85) assigns Health = Healthy at :74 without touching
// the GPU; DeviceMap.setEntry (device/device_map.go:101-111) stores it; the
// `health` channel ListAndWatch reads (plugin/plugin.go:181-186) has no
// writer.  gp_probe_device / gp_probe_all produce the verdict those sites
// consume.  There is deliberately no CPU fallback in this file: if CUDA is
// unavailable every call returns GP_ERR_CUDA / GP_ERR_NO_DEVICE.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>

#include "common.cuh"
#include "engine.h"

namespace gp {

using Stopwatch_t = std::chrono::steady_clock::time_point;

static thread_local char tls_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tls_error, sizeof(tls_error), fmt, ap);
  va_end(ap);
}

// The errors after which the context is gone: the runtime returns them again from every later call on
// the device, whatever the call (CUDA runtime API, "sticky" errors), so the only honest verdict is
// "lost" and the only recovery a reset or a fresh process.
bool cuda_error_is_sticky(cudaError_t e) {
  switch (e) {
    case cudaErrorIllegalAddress: case cudaErrorLaunchFailure: case cudaErrorHardwareStackError:
    case cudaErrorIllegalInstruction: case cudaErrorMisalignedAddress: case cudaErrorInvalidAddressSpace:
    case cudaErrorInvalidPc: case cudaErrorECCUncorrectable: case cudaErrorLaunchTimeout:
    case cudaErrorAssert: case cudaErrorNvlinkUncorrectable: case cudaErrorTensorMemoryLeak:
    case cudaErrorMpsClientTerminated: case cudaErrorExternalDevice: case cudaErrorDevicesUnavailable:
    case cudaErrorUnknown:
      return true;
    default:
      return false;
  }
}

static std::atomic<unsigned long long> g_poisoned{0};
unsigned long long poisoned_ordinals() { return g_poisoned.load(std::memory_order_acquire); }
void clear_poisoned_ordinal(int ordinal) {
  if (ordinal >= 0 && ordinal < 64) g_poisoned.fetch_and(~(1ull << ordinal), std::memory_order_acq_rel);
}

int cuda_fail(cudaError_t e, const char* what, int64_t* detail) {
  if (detail) *detail = (int64_t)e;
  if (cuda_error_is_sticky(e)) {
    int cur = -1;
    if (cudaGetDevice(&cur) == cudaSuccess && cur >= 0 && cur < 64)
      g_poisoned.fetch_or(1ull << cur, std::memory_order_acq_rel);
    set_error("%s: %s (%d) - sticky fault, the context on device %d is lost", what, cudaGetErrorString(e), (int)e, cur);
    (void)cudaGetLastError();
    return GP_ERR_DEVICE_LOST;
  }
  set_error("%s: %s (%d)", what, cudaGetErrorString(e), (int)e);
  (void)cudaGetLastError();
  return e == cudaErrorMemoryAllocation ? GP_ERR_NOMEM : GP_ERR_CUDA;
}

#define CU_OR_RETURN(call, detail)                                    \
  do {                                                                \
    cudaError_t e__ = (call);                                         \
    if (e__ != cudaSuccess) return ::gp::cuda_fail(e__, #call, detail); \
  } while (0)

static void format_uuid(const cudaUUID_t& u, char* out, size_t n) {
  const unsigned char* b = reinterpret_cast<const unsigned char*>(u.bytes);
  snprintf(out, n,
           "GPU-%02x%02x%02x%02x-%02x%02x-%02x%02x-%02x%02x-%02x%02x%02x%02x%02x%02x", b[0], b[1],
           b[2], b[3], b[4], b[5], b[6], b[7], b[8], b[9], b[10], b[11], b[12], b[13], b[14], b[15]);
}

static bool is_decimal(const char* s) {
  if (!*s) return false;
  for (; *s; ++s)
    if (*s < '0' || *s > '9') return false;
  return true;
}

// "GPU-<hex>" and "MIG-<hex>" name the same CUDA device when the hex part matches:
// under CUDA_VISIBLE_DEVICES=MIG-… the runtime reports the slice's own uuid bytes.
static const char* uuid_body(const char* s) {
  if (!strncmp(s, "GPU-", 4) || !strncmp(s, "MIG-", 4)) return s + 4;
  return s;
}

Device* find_device(Ctx* ctx, const char* uuid) {
  if (!ctx || !uuid) return nullptr;
  if (is_decimal(uuid)) {
    const long o = strtol(uuid, nullptr, 10);
    if (o >= 0 && o < (long)ctx->devs.size()) return ctx->devs[(size_t)o].get();
    return nullptr;
  }
  for (auto& d : ctx->devs)
    if (!strcasecmp(uuid_body(d->uuid), uuid_body(uuid))) return d.get();
  return nullptr;
}

int resolve_variant(const Ctx* ctx) {
  if (ctx->cfg.stream_variant == GP_STREAM_VEC || ctx->cfg.stream_variant == GP_STREAM_TMA)
    return (int)ctx->cfg.stream_variant;
  return GP_STREAM_TMA;  // AUTO: see DESIGN.md "Stream kernels" for the measured choice
}

static void device_release(Device* d) {
  release_arena_va(d);
  if (d->scratch) cudaFree(d->scratch);
  if (d->h_results) cudaFreeHost(d->h_results);
  for (auto& ev : d->events)
    if (ev) { cudaEventDestroy(ev); ev = nullptr; }
  if (d->stream) cudaStreamDestroy(d->stream);
  d->scratch = nullptr;
  d->h_results = nullptr;
  d->stream = nullptr;
  d->ready = false;
}

static int device_prepare_once(Device* d) {
  CU_OR_RETURN(cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking), nullptr);
  const size_t scratch_bytes =
      kVerifyScratchBytes + kMaxResults * sizeof(gp_pattern_result) + 256;
  CU_OR_RETURN(cudaMalloc(&d->scratch, scratch_bytes), nullptr);
  CU_OR_RETURN(cudaMemsetAsync(d->scratch, 0, scratch_bytes, d->stream), nullptr);
  CU_OR_RETURN(launch_verify_scratch_init(d->scratch, d->stream), nullptr);
  d->d_results = reinterpret_cast<gp_pattern_result*>(d->scratch + kVerifyScratchBytes);
  d->d_count = reinterpret_cast<unsigned long long*>(d->scratch + kVerifyScratchBytes +
                                                     kMaxResults * sizeof(gp_pattern_result));
  CU_OR_RETURN(cudaMallocHost(&d->h_results, kMaxResults * sizeof(gp_pattern_result) + 64), nullptr);
  d->h_count = reinterpret_cast<unsigned long long*>(d->h_results + kMaxResults);
  for (int i = 0; i < kEventPool; ++i) CU_OR_RETURN(cudaEventCreate(&d->events[i]), nullptr);
  CU_OR_RETURN(stream_kernels_configure(), nullptr);
  if (d->cc_major == 10) CU_OR_RETURN(gemm_kernels_configure(), nullptr);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), nullptr);
  return GP_OK;
}

int device_prepare(Ctx* ctx, Device* d) {
  (void)ctx;
  if (poisoned_ordinals() >> d->ordinal & 1) d->poisoned = true;
  if (d->poisoned) {
    set_error("device %d (%s) was lost to a sticky CUDA fault earlier in this process (cudaError %d): "
              "gp_device_reset it or grade it from a fresh process", d->ordinal, d->uuid, d->poison_code);
    return GP_ERR_DEVICE_LOST;
  }
  CU_OR_RETURN(cudaSetDevice(d->ordinal), nullptr);
  if (d->ready) return GP_OK;
  const int rc = device_prepare_once(d);
  if (rc != GP_OK) {  // nothing half-built survives a failed set-up
    device_release(d);
    return rc;
  }
  d->ready = true;
  return GP_OK;
}

static uint64_t cfg_hbm_bytes(const Ctx* c) { return c->cfg.hbm_bytes ? c->cfg.hbm_bytes : (16ull << 30); }
static uint64_t cfg_pattern_bytes(const Ctx* c) {
  return c->cfg.pattern_bytes ? c->cfg.pattern_bytes : cfg_hbm_bytes(c);
}
void summarize(const std::vector<float>& ms, gp_timing* t) {
  if (!t) return;
  memset(t, 0, sizeof(*t));
  if (ms.empty()) return;
  std::vector<float> s(ms);
  std::sort(s.begin(), s.end());
  double sum = 0;
  for (float v : s) sum += v;
  t->best_ms = s.front();
  t->median_ms = (s.size() & 1) ? s[s.size() / 2] : 0.5 * (s[s.size() / 2 - 1] + s[s.size() / 2]);
  t->mean_ms = sum / (double)s.size();
  t->launches = (uint32_t)s.size();
}

static void add_stat(Device* d, int kid, const std::vector<float>& ms, double algo_per_launch) {
  for (float v : ms) {
    d->stats[kid].launches += 1;
    d->stats[kid].total_ms += v;
    d->stats[kid].algo_bytes += algo_per_launch;
  }
}

static bool valid_kind_pass(int kind, int pass) {
  if (kind == GP_PAT_WALKING_ONES || kind == GP_PAT_WALKING_ZEROS) return pass >= 0 && pass < 64;
  if (kind == GP_PAT_ADDRESS) return pass == 0 || pass == 1;
  return false;
}

static double gbps(double bytes, double ms) { return ms > 0 ? bytes / (ms * 1e-3) / 1e9 : 0.0; }

// Holds the device lock with the device current and prepared.
struct Locked {
  std::unique_lock<std::mutex> lk;
  Device* d = nullptr;
  int rc = GP_OK;
  Locked(Ctx* ctx, const char* uuid) {
    d = find_device(ctx, uuid);
    if (!ctx) { rc = GP_ERR_INVALID; set_error("null context"); return; }
    if (!d) { rc = GP_ERR_NO_DEVICE; set_error("device '%s' is not visible to CUDA", uuid ? uuid : "(null)"); return; }
    lk = std::unique_lock<std::mutex>(d->mu);
    rc = device_prepare(ctx, d);
  }
};

// fill + verify for passes [pass_begin, pass_begin + count) of `kind`; device results land in
// d_results[slot0 ...]; per-launch times are appended to fill_ms / verify_ms.
static int pattern_run(Ctx* ctx, Device* d, uint64_t* words, int kind, int pass_begin, int count, uint64_t nwords,
                       int slot0, std::vector<float>* fill_ms, std::vector<float>* verify_ms,
                       int64_t* detail) {
  (void)ctx;
  if (count <= 0) return GP_OK;
  if (2 * count + 1 > kEventPool || slot0 + count > kMaxResults) return GP_ERR_TOO_LARGE;
  for (int j = 0; j < count; ++j) {
    CU_OR_RETURN(cudaEventRecord(d->events[2 * j], d->stream), detail);
    CU_OR_RETURN(launch_pattern_fill(words, nwords, kind, pass_begin + j, d->sm_count, d->stream), detail);
    CU_OR_RETURN(cudaEventRecord(d->events[2 * j + 1], d->stream), detail);
    CU_OR_RETURN(launch_pattern_verify(words, nwords, kind, pass_begin + j, d->scratch,
                                       d->d_results + slot0 + j, d->sm_count, d->stream), detail);
  }
  CU_OR_RETURN(cudaEventRecord(d->events[2 * count], d->stream), detail);
  CU_OR_RETURN(cudaMemcpyAsync(d->h_results + slot0, d->d_results + slot0,
                               (size_t)count * sizeof(gp_pattern_result), cudaMemcpyDeviceToHost,
                               d->stream), detail);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), detail);
  std::vector<float> f, v;
  for (int j = 0; j < count; ++j) {
    float a = 0, b = 0;
    CU_OR_RETURN(cudaEventElapsedTime(&a, d->events[2 * j], d->events[2 * j + 1]), detail);
    CU_OR_RETURN(cudaEventElapsedTime(&b, d->events[2 * j + 1], d->events[2 * j + 2]), detail);
    f.push_back(a);
    v.push_back(b);
  }
  add_stat(d, GP_K_FILL, f, (double)nwords * 8.0);
  add_stat(d, GP_K_VERIFY, v, (double)nwords * 8.0);
  if (fill_ms) fill_ms->insert(fill_ms->end(), f.begin(), f.end());
  if (verify_ms) verify_ms->insert(verify_ms->end(), v.begin(), v.end());
  return GP_OK;
}

static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
static inline uint64_t digest_step(uint64_t digest, const gp_pattern_result& r, uint32_t pass_no) {
  return digest + (r.sum ^ rotl64(r.xor_, 1) ^ rotl64(r.wsum, 2)) * (uint64_t)(pass_no + 1);
}

// `iters` timed launches of fn after `warmup` untimed ones
template <class F>
static int time_launches(Device* d, uint32_t warmup, uint32_t iters, F fn, std::vector<float>* ms,
                         int64_t* detail) {
  if (iters + 1 > (uint32_t)kEventPool) return GP_ERR_TOO_LARGE;
  for (uint32_t i = 0; i < warmup; ++i) CU_OR_RETURN(fn(), detail);
  for (uint32_t i = 0; i < iters; ++i) {
    CU_OR_RETURN(cudaEventRecord(d->events[i], d->stream), detail);
    CU_OR_RETURN(fn(), detail);
  }
  CU_OR_RETURN(cudaEventRecord(d->events[iters], d->stream), detail);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), detail);
  for (uint32_t i = 0; i < iters; ++i) {
    float t = 0;
    CU_OR_RETURN(cudaEventElapsedTime(&t, d->events[i], d->events[i + 1]), detail);
    ms->push_back(t);
  }
  return GP_OK;
}

// src and dst are `bytes` long each and need not be adjacent
static int copy_probe_run(Ctx* ctx, Device* d, uint64_t* src, uint8_t* dst, uint64_t bytes, uint32_t warmup,
                          uint32_t iters, gp_timing* t, uint64_t* checksum, uint32_t* ok, int64_t* detail) {
  if (bytes == 0 || (bytes & ((1u << 20) - 1))) { set_error("copy bytes must be a non-zero multiple of 1 MiB"); return GP_ERR_INVALID; }
  int rc = GP_OK;
  const int variant = resolve_variant(ctx);
  const uint64_t nwords = bytes / 8;
  CU_OR_RETURN(launch_pattern_fill(src, nwords, GP_PAT_ADDRESS, 0, d->sm_count, d->stream), detail);
  CU_OR_RETURN(cudaMemsetAsync(dst, 0xA5, bytes, d->stream), detail);
  std::vector<float> ms;
  rc = time_launches(d, warmup, iters, [&] { return launch_copy(dst, src, bytes, variant, d->sm_count, d->stream); }, &ms, detail);
  if (rc) return rc;
  add_stat(d, GP_K_COPY, ms, 2.0 * (double)bytes);
  summarize(ms, t);
  // dst must now verify as the address pattern, word for word
  CU_OR_RETURN(launch_pattern_verify(reinterpret_cast<uint64_t*>(dst), nwords, GP_PAT_ADDRESS, 0,
                                     d->scratch, d->d_results, d->sm_count, d->stream), detail);
  CU_OR_RETURN(cudaMemcpyAsync(d->h_results, d->d_results, sizeof(gp_pattern_result),
                               cudaMemcpyDeviceToHost, d->stream), detail);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), detail);
  if (checksum) *checksum = d->h_results[0].sum;
  if (ok) *ok = d->h_results[0].mismatches == 0;
  return GP_OK;
}

static int copy_probe_locked(Ctx* ctx, Device* d, uint64_t bytes, uint32_t warmup, uint32_t iters,
                             gp_timing* t, uint64_t* checksum, uint32_t* ok, int64_t* detail) {
  if (bytes == 0 || (bytes & ((1u << 20) - 1))) { set_error("copy bytes must be a non-zero multiple of 1 MiB"); return GP_ERR_INVALID; }
  int rc = ensure_arena(ctx, d, 2 * bytes);
  if (rc) return rc;
  return copy_probe_run(ctx, d, reinterpret_cast<uint64_t*>(d->arena), d->arena + bytes, bytes, warmup, iters, t, checksum, ok, detail);
}

// a, b, c are `bytes` long each and need not be adjacent
static int triad_probe_run(Ctx* ctx, Device* d, float* a, float* b, float* c, uint64_t bytes, uint32_t warmup,
                           uint32_t iters, gp_timing* t, uint64_t* bad, int64_t* detail, uint64_t* sum_xor) {
  if (bytes == 0 || (bytes & ((1u << 20) - 1))) { set_error("triad bytes must be a non-zero multiple of 1 MiB"); return GP_ERR_INVALID; }
  int rc = GP_OK;
  d->triad_operands_bytes = 0;  // the step layout's operands are about to be overwritten
  const int variant = resolve_variant(ctx);
  const uint64_t seed = ctx->cfg.seed ? ctx->cfg.seed : 0x5EEDull;
  const uint64_t n = bytes / 4;
  const float s = 3.0f;
  CU_OR_RETURN(launch_gen_f32(b, n, seed ^ kSeedTriadBXor, d->sm_count, d->stream), detail);
  CU_OR_RETURN(launch_gen_f32(c, n, seed ^ kSeedTriadCXor, d->sm_count, d->stream), detail);
  CU_OR_RETURN(cudaMemsetAsync(a, 0xFF, bytes, d->stream), detail);
  std::vector<float> ms;
  rc = time_launches(d, warmup, iters, [&] { return launch_triad(a, b, c, s, bytes, variant, d->sm_count, d->stream); }, &ms, detail);
  if (rc) return rc;
  add_stat(d, GP_K_TRIAD, ms, 3.0 * (double)bytes);
  summarize(ms, t);
  CU_OR_RETURN(cudaMemsetAsync(d->d_count, 0, sizeof(unsigned long long), d->stream), detail);
  CU_OR_RETURN(launch_triad_check(a, n, seed, s, d->d_count, d->sm_count, d->stream), detail);
  CU_OR_RETURN(cudaMemcpyAsync(d->h_count, d->d_count, sizeof(unsigned long long),
                               cudaMemcpyDeviceToHost, d->stream), detail);
  // the result's checksum, for the known-answer compare (the check above is the GPU agreeing with itself)
  CU_OR_RETURN(launch_pattern_verify(reinterpret_cast<const uint64_t*>(a), bytes / 8, kPatNone, 0, d->scratch,
                                     d->d_results, d->sm_count, d->stream), detail);
  CU_OR_RETURN(cudaMemcpyAsync(d->h_results, d->d_results, sizeof(gp_pattern_result), cudaMemcpyDeviceToHost,
                               d->stream), detail);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), detail);
  if (bad) *bad = *d->h_count;
  if (sum_xor) { sum_xor[0] = d->h_results[0].sum; sum_xor[1] = d->h_results[0].xor_; }
  return GP_OK;
}

static int triad_probe_locked(Ctx* ctx, Device* d, uint64_t bytes, uint32_t warmup, uint32_t iters,
                              gp_timing* t, uint64_t* bad, int64_t* detail, uint64_t* sum_xor = nullptr) {
  if (bytes == 0 || (bytes & ((1u << 20) - 1))) { set_error("triad bytes must be a non-zero multiple of 1 MiB"); return GP_ERR_INVALID; }
  int rc = ensure_arena(ctx, d, 3 * bytes);
  if (rc) return rc;
  return triad_probe_run(ctx, d, reinterpret_cast<float*>(d->arena), reinterpret_cast<float*>(d->arena + bytes),
                         reinterpret_cast<float*>(d->arena + 2 * bytes), bytes, warmup, iters, t, bad, detail, sum_xor);
}

static int probe_device_impl(Ctx* ctx, const char* uuid, uint32_t mask, gp_device_result* out) {
  if (!out) { set_error("null result"); return GP_ERR_INVALID; }
  const auto t0 = std::chrono::steady_clock::now();
  memset(out, 0, sizeof(*out));
  out->struct_size = (uint32_t)sizeof(*out);
  out->abi_version = GP_ABI_VERSION;
  out->first_bad_kind = -1;
  out->first_bad_pass = -1;
  out->first_bad_index = ~0ull;
  Locked L(ctx, uuid);
  if (L.rc) { out->error = L.rc; return L.rc; }
  Device* d = L.d;
  d->alloc_s = 0;
  d->free_s = 0;
  out->arena_reused = d->arena != nullptr;
  snprintf(out->uuid, sizeof(out->uuid), "%s", d->uuid);
  out->ordinal = d->ordinal;
  out->sm_count = d->sm_count;
  out->total_mem = d->total_mem;
  out->cc_major = d->cc_major;
  out->cc_minor = d->cc_minor;
  const uint32_t warmup = ctx->cfg.warmup ? ctx->cfg.warmup : 2;
  const uint32_t iters = ctx->cfg.iters ? ctx->cfg.iters : 10;
  const uint64_t hbm = cfg_hbm_bytes(ctx);
  uint64_t pat = cfg_pattern_bytes(ctx);
  const bool pat_all = pat == GP_PATTERN_ALL_FREE && (mask & GP_PROBE_PATTERN);
  if (pat == GP_PATTERN_ALL_FREE && !pat_all) pat = hbm;  // no pattern probe asked for: nothing to size
  if (pat_all) {
    // whole-memory grading: everything the driver will still give us, minus a margin, in 1 MiB units
    size_t free_b = 0, total_b = 0;
    free_arena(d);
    cudaError_t e = cudaMemGetInfo(&free_b, &total_b);
    if (e != cudaSuccess) { out->error = cuda_fail(e, "cudaMemGetInfo", &out->detail); return out->error; }
    const uint64_t margin = 1ull << 30;
    pat = free_b > 2 * margin ? ((uint64_t)free_b - margin) & ~((1ull << 20) - 1) : 0;
  }
  out->hbm_bytes = hbm;
  out->pattern_bytes = pat;
  int rc = GP_OK;
  // ---- where the arrays live -------------------------------------------------------------------------------
  //   x : pattern words, copy source, triad result      y : copy destination, triad operand b      z : triad operand c
  // Kept arena (or whole-memory grading): x, y, z are consecutive ranges of the one arena.  Cold call (keep_arena = 0,
  // cudaMalloc backend): three separate slabs, allocated IN ORDER on a provider thread while the probes that do not
  // need them yet are already running — the GEMM's checks cover the first slab, the pattern probe's 66 passes over x
  // cover the other two — and handed to the reaper thread when the call returns.  Nothing requires x, y and z to be
  // adjacent, and the driver serialises every GPU's allocations, so a call that waits for one 48 GiB allocation
  // waits three times longer than one that waits for its first 16 GiB.
  struct Slabs {
    std::mutex mu;
    std::condition_variable cv;
    int ready = 0;                 // slabs [0, ready) are allocated
    cudaError_t err = cudaSuccess;
    void* p[3] = {nullptr, nullptr, nullptr};
    size_t bytes[3] = {0, 0, 0};
    std::thread provider;
  } slabs;
  const uint32_t gemm_mnk = ctx->cfg.gemm_mnk ? ctx->cfg.gemm_mnk : 8192;
  const uint32_t gemm_samples = ctx->cfg.gemm_samples ? ctx->cfg.gemm_samples : 4096;
  const uint64_t gemm_need = (mask & GP_PROBE_GEMM) ? gemm_probe_bytes(ctx, gemm_mnk, gemm_samples) : 0;
  uint64_t hbm_need = 0;
  if (mask & GP_PROBE_COPY) hbm_need = std::max<uint64_t>(hbm_need, 2 * hbm);
  if (mask & GP_PROBE_TRIAD) hbm_need = std::max<uint64_t>(hbm_need, 3 * hbm);
  if (mask & GP_PROBE_PATTERN) hbm_need = std::max<uint64_t>(hbm_need, pat);
  const bool use_slabs = !ctx->cfg.keep_arena && !pat_all && hbm_need > 0 && !d->arena &&
                         ctx->cfg.arena_backend != GP_ARENA_VMM;
  if (use_slabs) {
    if (mask & (GP_PROBE_COPY | GP_PROBE_TRIAD)) slabs.bytes[0] = hbm;
    if (mask & GP_PROBE_PATTERN) slabs.bytes[0] = std::max<uint64_t>(slabs.bytes[0], pat);
    if (mask & (GP_PROBE_COPY | GP_PROBE_TRIAD)) slabs.bytes[1] = hbm;
    if (mask & GP_PROBE_TRIAD) slabs.bytes[2] = hbm;
  }
  uint64_t slab_total = 0;
  void* gemm_ws = nullptr;  // the GEMM's own buffer in slab mode; released with the slabs (a cudaFree here would
                            // queue behind every other GPU's slab allocation: the driver serialises them)

  auto finish = [&] {
    if (slabs.provider.joinable()) slabs.provider.join();
    if (use_slabs) {
      std::vector<void*> ptrs;
      for (void* q : slabs.p) if (q) ptrs.push_back(q);
      if (gemm_ws) ptrs.push_back(gemm_ws);
      out->arena_bytes = slab_total;
      out->arena_backend = ptrs.empty() ? 0 : GP_ARENA_MALLOC;
      out->free_deferred = !ptrs.empty();
      if (d->arena) free_arena(d);     // (only the not-sm_100 GEMM path puts anything there in this mode)
      free_buffers_deferred(d, ptrs);  // the results are complete: the memory goes back on the reaper thread
    } else {
      out->arena_bytes = d->arena_bytes;
      out->arena_backend = d->arena ? (d->arena_is_vmm ? GP_ARENA_VMM : GP_ARENA_MALLOC) : 0;
      if (!ctx->cfg.keep_arena) { out->free_deferred = d->arena != nullptr; free_arena_deferred(d); }
    }
    out->duration_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    out->alloc_s = d->alloc_s;
    out->free_s = d->free_s;
    out->kernels_s = out->duration_s - out->alloc_s - out->free_s;
  };
  auto fail = [&](int code) {
    out->error = code;
    out->healthy = 0;
    if (code == GP_ERR_DEVICE_LOST) { d->poisoned = true; d->poison_code = (int)out->detail; }
    finish();
    return code;
  };

  if (use_slabs) {
    slab_total = slabs.bytes[0] + slabs.bytes[1] + slabs.bytes[2];
    if (slab_total > cfg_arena_cap(ctx)) {
      set_error("arena request %llu exceeds cap %llu", (unsigned long long)slab_total, (unsigned long long)cfg_arena_cap(ctx));
      return fail(GP_ERR_NOMEM);
    }
  }
  const std::function<void()> start_provider = [&] {
    if (!use_slabs || slabs.provider.joinable()) return;
    slabs.provider = std::thread([&slabs, d] {
      cudaSetDevice(d->ordinal);
      arena_join_reaper(d);  // the previous call's memory must be back before this one asks for as much again
      for (int i = 0; i < 3; ++i) {
        cudaError_t e = cudaSuccess;
        void* q = nullptr;
        if (slabs.bytes[i]) e = cudaMalloc(&q, slabs.bytes[i]);
        std::lock_guard<std::mutex> g(slabs.mu);
        slabs.p[i] = q;
        if (e != cudaSuccess) { slabs.err = e; (void)cudaGetLastError(); slabs.ready = 3; slabs.cv.notify_all(); return; }
        slabs.ready = i + 1;
        slabs.cv.notify_all();
      }
    });
  };
  // block until slab i exists; the wait is the only part of the allocation this call pays for
  auto wait_slab = [&](int i) -> int {
    const auto w0 = std::chrono::steady_clock::now();
    std::unique_lock<std::mutex> lk(slabs.mu);
    slabs.cv.wait(lk, [&] { return slabs.ready > i; });
    d->alloc_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
    if (slabs.err != cudaSuccess && !slabs.p[i]) return cuda_fail(slabs.err, "cudaMalloc(probe arrays)", &out->detail);
    return GP_OK;
  };
  auto grow = [&](uint64_t bytes) -> int {
    if (bytes > d->arena_bytes && (bytes <= cfg_arena_cap(ctx) || pat_all)) return ensure_arena(ctx, d, bytes, /*ignore_cap=*/pat_all);
    if (bytes > d->arena_bytes) { set_error("arena request %llu exceeds cap %llu", (unsigned long long)bytes, (unsigned long long)cfg_arena_cap(ctx)); return GP_ERR_NOMEM; }
    return GP_OK;
  };
  uint8_t *bx = nullptr, *by = nullptr, *bz = nullptr;
  auto arrays = [&](int upto) -> int {  // make x (0), y (1), z (2) usable
    if (use_slabs) {
      int r = wait_slab(upto);
      if (r) return r;
      bx = reinterpret_cast<uint8_t*>(slabs.p[0]); by = reinterpret_cast<uint8_t*>(slabs.p[1]); bz = reinterpret_cast<uint8_t*>(slabs.p[2]);
      return GP_OK;
    }
    bx = d->arena; by = d->arena + hbm; bz = d->arena + 2 * hbm;
    return GP_OK;
  };

  auto run_gemm = [&](uint8_t* workspace, const std::function<void()>* after_timed) -> int {
    const uint64_t seed = ctx->cfg.seed ? ctx->cfg.seed : 0x5EEDull;
    // The GEMM is the first thing this call runs on a GPU that may have sat idle: give the clocks a few
    // milliseconds of tensor work to settle before anything is timed (8 launches at 8192^3 = 5.5 ms).
    const uint32_t gemm_warmup = std::max<uint32_t>(warmup, 8);
    int r = gemm_probe_locked(ctx, d, gemm_mnk, seed, gemm_warmup, iters, gemm_samples, &out->gemm, workspace, after_timed);
    if (r) return r;
    out->probed_mask |= GP_PROBE_GEMM;
    if (out->gemm.samples_bad ||
        (ctx->cfg.min_gemm_tflops > 0 && out->gemm.tflops_median < ctx->cfg.min_gemm_tflops))
      out->failed_mask |= GP_PROBE_GEMM;
    return GP_OK;
  };

  // ---- GEMM first: on a cool GPU, and (cold call) with the allocator busy behind its checks --------------
  if (mask & GP_PROBE_GEMM) {
    if (use_slabs && d->cc_major == 10) {
      // a small buffer of its own (a millisecond to allocate).  The provider thread starts once the timed launches
      // are over: allocating DURING them was tried first and cost the GEMM 9 % of its median.
      const auto a0 = std::chrono::steady_clock::now();
      cudaError_t e = cudaMalloc(&gemm_ws, gemm_need);
      d->alloc_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - a0).count();
      if (e != cudaSuccess) { gemm_ws = nullptr; return fail(cuda_fail(e, "cudaMalloc(gemm workspace)", &out->detail)); }
      rc = run_gemm(reinterpret_cast<uint8_t*>(gemm_ws), &start_provider);
      start_provider();  // (if the GEMM failed before its timed launches ended)
      if (rc) return fail(rc);
    } else {
      // (grow before the provider starts: both would otherwise wait for the same reaper thread)
      rc = grow(use_slabs ? gemm_need : std::max(hbm_need, gemm_need));  // use_slabs here = not sm_100: run_gemm reports it
      if (rc) return fail(rc);
      start_provider();
      rc = run_gemm(nullptr, nullptr);
      if (rc) return fail(rc);
    }
  } else {
    start_provider();
    if (!use_slabs && hbm_need) { rc = grow(hbm_need); if (rc) return fail(rc); }
  }

  if (mask & GP_PROBE_PATTERN) {
    if (pat == 0 || (pat % 512)) { set_error("pattern_bytes must be a multiple of 512"); return fail(GP_ERR_INVALID); }
    rc = arrays(0);
    if (rc) return fail(rc);
    d->triad_operands_bytes = 0;
    const uint64_t nwords = pat / 8;
    int wo = ctx->cfg.wo_passes ? (int)ctx->cfg.wo_passes : 64;
    if (wo > 64) wo = 64;
    std::vector<float> fms, vms;
    uint64_t* words = reinterpret_cast<uint64_t*>(bx);
    rc = pattern_run(ctx, d, words, GP_PAT_WALKING_ONES, 0, wo, nwords, 0, &fms, &vms, &out->detail);
    if (rc) return fail(rc);
    rc = pattern_run(ctx, d, words, GP_PAT_ADDRESS, 0, 2, nwords, wo, &fms, &vms, &out->detail);
    if (rc) return fail(rc);
    out->probed_mask |= GP_PROBE_PATTERN;
    summarize(fms, &out->fill_t);
    summarize(vms, &out->verify_t);
    out->fill_gbps_best = gbps((double)pat, out->fill_t.best_ms);
    out->fill_gbps_median = gbps((double)pat, out->fill_t.median_ms);
    out->verify_gbps_best = gbps((double)pat, out->verify_t.best_ms);
    out->verify_gbps_median = gbps((double)pat, out->verify_t.median_ms);
    out->pattern_passes = (uint32_t)(wo + 2);
    uint64_t digest = 0;
    for (int j = 0; j < wo + 2; ++j) {
      const gp_pattern_result& r = d->h_results[j];
      digest = digest_step(digest, r, (uint32_t)j);
      if (r.mismatches) {
        out->pattern_bad_passes++;
        out->pattern_mismatches += r.mismatches;
        if (out->first_bad_kind < 0) {
          out->first_bad_kind = j < wo ? GP_PAT_WALKING_ONES : GP_PAT_ADDRESS;
          out->first_bad_pass = j < wo ? j : j - wo;
          out->first_bad_index = r.first_bad_index;
          out->first_bad_xor = r.first_bad_xor;
        }
      }
    }
    out->pattern_digest = digest;
    if (out->pattern_bad_passes) out->failed_mask |= GP_PROBE_PATTERN;
  }
  if (mask & GP_PROBE_COPY) {
    rc = arrays(1);
    if (rc) return fail(rc);
    d->triad_operands_bytes = 0;
    rc = copy_probe_run(ctx, d, reinterpret_cast<uint64_t*>(bx), by, hbm, warmup, iters, &out->copy_t, &out->copy_checksum,
                        &out->copy_ok, &out->detail);
    if (rc) return fail(rc);
    out->probed_mask |= GP_PROBE_COPY;
    out->copy_gbps_best = gbps(2.0 * hbm, out->copy_t.best_ms);
    out->copy_gbps_median = gbps(2.0 * hbm, out->copy_t.median_ms);
    if (!out->copy_ok || (ctx->cfg.min_copy_gbps > 0 && out->copy_gbps_median < ctx->cfg.min_copy_gbps))
      out->failed_mask |= GP_PROBE_COPY;
  }
  if (mask & GP_PROBE_TRIAD) {
    rc = arrays(2);
    if (rc) return fail(rc);
    uint64_t bad = 0;
    uint64_t sx[2] = {0, 0};
    rc = triad_probe_run(ctx, d, reinterpret_cast<float*>(bx), reinterpret_cast<float*>(by), reinterpret_cast<float*>(bz), hbm,
                         warmup, iters, &out->triad_t, &bad, &out->detail, sx);
    if (rc) return fail(rc);
    out->probed_mask |= GP_PROBE_TRIAD;
    out->triad_ok = bad == 0;
    out->triad_sum = sx[0];
    out->triad_xor = sx[1];
    if (ctx->cfg.triad_expect_sum || ctx->cfg.triad_expect_xor) {
      out->triad_known_answer = (sx[0] == ctx->cfg.triad_expect_sum && sx[1] == ctx->cfg.triad_expect_xor) ? 1 : 2;
      if (out->triad_known_answer == 2) out->failed_mask |= GP_PROBE_TRIAD;
    }
    out->triad_gbps_best = gbps(3.0 * hbm, out->triad_t.best_ms);
    out->triad_gbps_median = gbps(3.0 * hbm, out->triad_t.median_ms);
    if (!out->triad_ok || (ctx->cfg.min_triad_gbps > 0 && out->triad_gbps_median < ctx->cfg.min_triad_gbps))
      out->failed_mask |= GP_PROBE_TRIAD;
  }
  out->healthy = out->failed_mask == 0;
  finish();
  return GP_OK;
}

}  // namespace gp

using namespace gp;

// ============================================================= C ABI ========
extern "C" {

uint32_t gp_abi_version(void) { return GP_ABI_VERSION; }

size_t gp_sizeof(int which) {
  switch (which) {
    case 0: return sizeof(gp_config);
    case 1: return sizeof(gp_pattern_result);
    case 2: return sizeof(gp_timing);
    case 3: return sizeof(gp_gemm_result);
    case 4: return sizeof(gp_device_result);
    case 5: return sizeof(gp_kernel_stat);
    case 6: return sizeof(gp_nvml_info);
    case 7: return sizeof(gp_fabric_result);
    default: return 0;
  }
}

const char* gp_strerror(int err) {
  switch (err) {
    case GP_OK: return "ok";
    case GP_ERR_INVALID: return "invalid argument";
    case GP_ERR_NO_DEVICE: return "device not visible to CUDA";
    case GP_ERR_CUDA: return "CUDA error";
    case GP_ERR_NCCL: return "NCCL error";
    case GP_ERR_NOMEM: return "out of device memory or arena cap exceeded";
    case GP_ERR_UNSUPPORTED: return "unsupported on this device";
    case GP_ERR_STRUCT_SIZE: return "struct size / ABI version mismatch";
    case GP_ERR_TOO_LARGE: return "request exceeds a fixed limit";
    case GP_ERR_DEVICE_LOST: return "device lost to a sticky CUDA fault";
    default: return "unknown error";
  }
}

const char* gp_last_error_detail(void) { return tls_error; }

int gp_init(const gp_config* cfg, gp_ctx** out) {
  if (!out) { set_error("null out"); return GP_ERR_INVALID; }
  *out = nullptr;
  gp_config c;
  memset(&c, 0, sizeof(c));
  if (cfg) {
    if (cfg->struct_size != sizeof(gp_config) || cfg->abi_version != GP_ABI_VERSION) {
      set_error("gp_config struct_size %u / abi %u, library expects %zu / %u", cfg->struct_size,
                cfg->abi_version, sizeof(gp_config), GP_ABI_VERSION);
      return GP_ERR_STRUCT_SIZE;
    }
    c = *cfg;
  }
  c.struct_size = sizeof(gp_config);
  c.abi_version = GP_ABI_VERSION;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount", nullptr);
  if (n == 0) { set_error("no CUDA device"); return GP_ERR_NO_DEVICE; }
  gp_ctx* ctx = new gp_ctx();
  ctx->cfg = c;
  for (int i = 0; i < n; ++i) {
    cudaDeviceProp p;
    e = cudaGetDeviceProperties(&p, i);
    if (e != cudaSuccess) { delete ctx; return cuda_fail(e, "cudaGetDeviceProperties", nullptr); }
    auto d = std::make_unique<Device>();
    d->ordinal = i;
    format_uuid(p.uuid, d->uuid, sizeof(d->uuid));
    d->sm_count = p.multiProcessorCount;
    if (c.sm_limit && (int)c.sm_limit < d->sm_count) d->sm_count = (int)c.sm_limit;
    d->cc_major = p.major;
    d->cc_minor = p.minor;
    d->total_mem = p.totalGlobalMem;
    ctx->devs.push_back(std::move(d));
  }
  *out = ctx;
  return GP_OK;
}

void gp_shutdown(gp_ctx* ctx) {
  if (!ctx) return;
  nccl_teardown(ctx);
  for (auto& d : ctx->devs) {
    std::lock_guard<std::mutex> g(d->mu);
    arena_join_reaper(d.get());
    if (!d->ready && !d->arena && !d->va_base) continue;
    cudaSetDevice(d->ordinal);
    if (d->stream && !d->poisoned) cudaStreamSynchronize(d->stream);
    device_release(d.get());
  }
  delete ctx;
}

int gp_trim(gp_ctx* ctx) {
  if (!ctx) return GP_ERR_INVALID;
  for (auto& d : ctx->devs) {
    std::lock_guard<std::mutex> g(d->mu);
    if (d->stream) { cudaSetDevice(d->ordinal); cudaStreamSynchronize(d->stream); }
    free_arena(d.get());
  }
  return GP_OK;
}

int gp_trim_device(gp_ctx* ctx, const char* uuid) {
  Device* d = find_device(ctx, uuid);
  if (!d) { set_error("device '%s' is not visible to CUDA", uuid ? uuid : "(null)"); return ctx ? GP_ERR_NO_DEVICE : GP_ERR_INVALID; }
  std::lock_guard<std::mutex> g(d->mu);
  if (d->stream && !d->poisoned) { cudaSetDevice(d->ordinal); cudaStreamSynchronize(d->stream); }
  free_arena(d);
  return GP_OK;
}

int gp_device_poisoned(gp_ctx* ctx, const char* uuid) {
  Device* d = find_device(ctx, uuid);
  if (!d) return ctx ? GP_ERR_NO_DEVICE : GP_ERR_INVALID;
  std::lock_guard<std::mutex> g(d->mu);
  return (d->poisoned || (poisoned_ordinals() >> d->ordinal & 1)) ? 1 : 0;
}

int gp_device_reset(gp_ctx* ctx, const char* uuid) {
  Device* d = find_device(ctx, uuid);
  if (!d) { set_error("device '%s' is not visible to CUDA", uuid ? uuid : "(null)"); return ctx ? GP_ERR_NO_DEVICE : GP_ERR_INVALID; }
  std::lock_guard<std::mutex> multi(ctx->multi_mu);  // no cross-device probe while a context goes away
  // Everything that points INTO the dying context from outside has to go first: communicators, the peer
  // mappings the other devices hold on this one, and their arenas' access grants for it.  (Resetting with those
  // in place crashed inside the driver later, in cuEventDestroy at shutdown: seen on 2 and 8 GPUs.)
  nccl_teardown(ctx);
  for (auto& o : ctx->devs) {
    if (o.get() == d) continue;
    std::lock_guard<std::mutex> g(o->mu);
    if (!o->ready && !o->arena) continue;
    arena_join_reaper(o.get());
    if (cudaSetDevice(o->ordinal) != cudaSuccess) { (void)cudaGetLastError(); continue; }
    if (o->stream) cudaStreamSynchronize(o->stream);
    if (o->peer_mask >> d->ordinal & 1) free_arena(o.get());  // its mapping names the dying device as an accessor
    (void)cudaDeviceDisablePeerAccess(d->ordinal);             // "not enabled" is the common, harmless answer
    (void)cudaGetLastError();
  }
  std::lock_guard<std::mutex> g(d->mu);
  cudaError_t e = cudaSetDevice(d->ordinal);
  (void)e;
  for (auto& o : ctx->devs)
    if (o.get() != d) (void)cudaDeviceDisablePeerAccess(o->ordinal);
  (void)cudaGetLastError();
  // every handle below belongs to the context that is about to die; failures while dropping them are expected
  device_release(d);
  (void)cudaGetLastError();
  noalloc_forget_device(d->ordinal);
  cudaDeviceSynchronize();
  (void)cudaGetLastError();
  e = cudaDeviceReset();
  if (e != cudaSuccess) {
    set_error("cudaDeviceReset: %s (%d)", cudaGetErrorString(e), (int)e);
    (void)cudaGetLastError();
    return GP_ERR_DEVICE_LOST;
  }
  d->poisoned = false;
  d->poison_code = 0;
  clear_poisoned_ordinal(d->ordinal);
  memset(d->stats, 0, sizeof(d->stats));
  return GP_OK;
}

int gp_inject_cuda_error(gp_ctx* ctx, const char* uuid, int cuda_error) {
  Device* d = find_device(ctx, uuid);
  if (!d) return ctx ? GP_ERR_NO_DEVICE : GP_ERR_INVALID;
  std::lock_guard<std::mutex> g(d->mu);
  cudaError_t e = cudaSetDevice(d->ordinal);
  if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice", nullptr);
  int64_t detail = 0;
  const int rc = cuda_fail((cudaError_t)cuda_error, "injected fault (gp_inject_cuda_error)", &detail);
  if (rc == GP_ERR_DEVICE_LOST) { d->poisoned = true; d->poison_code = cuda_error; }
  return rc;
}

int gp_device_count(gp_ctx* ctx) { return ctx ? (int)ctx->devs.size() : GP_ERR_INVALID; }

int gp_device_uuid(gp_ctx* ctx, int ordinal, char* buf, size_t buflen) {
  if (!ctx || !buf || buflen == 0) return GP_ERR_INVALID;
  if (ordinal < 0 || ordinal >= (int)ctx->devs.size()) return GP_ERR_NO_DEVICE;
  snprintf(buf, buflen, "%s", ctx->devs[(size_t)ordinal]->uuid);
  return GP_OK;
}

int gp_probe_device(gp_ctx* ctx, const char* uuid, uint32_t which_mask, gp_device_result* out) {
  if (!ctx) { set_error("null context"); return GP_ERR_INVALID; }
  return probe_device_impl(ctx, uuid, which_mask, out);
}

int gp_probe_all(gp_ctx* ctx, int ngpus, uint32_t which_mask, gp_device_result* out, double* wall_s) {
  if (!ctx || !out || ngpus < 1) { set_error("bad arguments"); return GP_ERR_INVALID; }
  if (ngpus > (int)ctx->devs.size()) { set_error("ngpus %d > visible %zu", ngpus, ctx->devs.size()); return GP_ERR_NO_DEVICE; }
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<int> rcs((size_t)ngpus, 0);
  std::vector<std::thread> th;
  for (int i = 0; i < ngpus; ++i)
    th.emplace_back([&, i] { rcs[(size_t)i] = probe_device_impl(ctx, ctx->devs[(size_t)i]->uuid, which_mask, &out[i]); });
  for (auto& t : th) t.join();
  if (wall_s) *wall_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  for (int rc : rcs)
    if (rc) return rc;
  return GP_OK;
}

int gp_pattern_fill(gp_ctx* ctx, const char* uuid, int kind, int pass, uint64_t nwords) {
  if (!valid_kind_pass(kind, pass) || (nwords % 64)) { set_error("bad kind/pass or nwords %% 64 != 0"); return GP_ERR_INVALID; }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  int rc = ensure_arena(ctx, L.d, nwords * 8);
  if (rc) return rc;
  L.d->triad_operands_bytes = 0;
  CU_OR_RETURN(launch_pattern_fill(reinterpret_cast<uint64_t*>(L.d->arena), nwords, kind, pass, L.d->sm_count, L.d->stream), nullptr);
  CU_OR_RETURN(cudaStreamSynchronize(L.d->stream), nullptr);
  return GP_OK;
}

int gp_pattern_verify(gp_ctx* ctx, const char* uuid, int kind, int pass, uint64_t nwords,
                      gp_pattern_result* out) {
  if (!out || !valid_kind_pass(kind, pass) || (nwords % 64)) { set_error("bad kind/pass, null out or nwords %% 64 != 0"); return GP_ERR_INVALID; }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  Device* d = L.d;
  if (nwords * 8 > d->arena_bytes) { set_error("verify range exceeds the arena (fill first)"); return GP_ERR_INVALID; }
  CU_OR_RETURN(launch_pattern_verify(reinterpret_cast<uint64_t*>(d->arena), nwords, kind, pass, d->scratch, d->d_results, d->sm_count, d->stream), nullptr);
  CU_OR_RETURN(cudaMemcpyAsync(d->h_results, d->d_results, sizeof(gp_pattern_result), cudaMemcpyDeviceToHost, d->stream), nullptr);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), nullptr);
  *out = d->h_results[0];
  return GP_OK;
}

int gp_pattern_host(gp_ctx* ctx, const char* uuid, uint64_t* host_words, uint64_t nwords, int kind,
                    int pass, int fill, gp_pattern_result* out) {
  if (!out || !valid_kind_pass(kind, pass) || (nwords % 64) || (!host_words && nwords && !fill)) {
    set_error("bad kind/pass, null pointer or nwords %% 64 != 0");
    return GP_ERR_INVALID;
  }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  Device* d = L.d;
  int rc = ensure_arena(ctx, d, std::max<uint64_t>(nwords * 8, 512));
  if (rc) return rc;
  d->triad_operands_bytes = 0;
  uint64_t* words = reinterpret_cast<uint64_t*>(d->arena);
  if (fill) {
    CU_OR_RETURN(launch_pattern_fill(words, nwords, kind, pass, d->sm_count, d->stream), nullptr);
  } else if (nwords) {
    CU_OR_RETURN(cudaMemcpyAsync(words, host_words, nwords * 8, cudaMemcpyHostToDevice, d->stream), nullptr);
  }
  CU_OR_RETURN(launch_pattern_verify(words, nwords, kind, pass, d->scratch, d->d_results, d->sm_count, d->stream), nullptr);
  CU_OR_RETURN(cudaMemcpyAsync(d->h_results, d->d_results, sizeof(gp_pattern_result), cudaMemcpyDeviceToHost, d->stream), nullptr);
  if (fill && host_words && nwords)
    CU_OR_RETURN(cudaMemcpyAsync(host_words, words, nwords * 8, cudaMemcpyDeviceToHost, d->stream), nullptr);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), nullptr);
  *out = d->h_results[0];
  return GP_OK;
}

int gp_pattern_probe(gp_ctx* ctx, const char* uuid, int kind, int pass_begin, int pass_count,
                     uint64_t nwords, gp_pattern_result* results, gp_timing* fill_t,
                     gp_timing* verify_t) {
  if (pass_count < 1 || !valid_kind_pass(kind, pass_begin) ||
      !valid_kind_pass(kind, pass_begin + pass_count - 1) || (nwords % 64) || nwords == 0) {
    set_error("bad kind/pass range or nwords");
    return GP_ERR_INVALID;
  }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  Device* d = L.d;
  int rc = ensure_arena(ctx, d, nwords * 8);
  if (rc) return rc;
  d->triad_operands_bytes = 0;
  std::vector<float> f, v;
  rc = pattern_run(ctx, d, reinterpret_cast<uint64_t*>(d->arena), kind, pass_begin, pass_count, nwords, 0, &f, &v, nullptr);
  if (rc) return rc;
  if (results) memcpy(results, d->h_results, (size_t)pass_count * sizeof(gp_pattern_result));
  summarize(f, fill_t);
  summarize(v, verify_t);
  return GP_OK;
}

int gp_stream_probe(gp_ctx* ctx, const char* uuid, uint32_t which, uint64_t bytes, uint32_t warmup,
                    uint32_t iters, gp_timing* t, uint64_t* checksum) {
  if (which != GP_PROBE_COPY && which != GP_PROBE_TRIAD) { set_error("which must be COPY or TRIAD"); return GP_ERR_INVALID; }
  if (iters == 0) { set_error("iters must be > 0"); return GP_ERR_INVALID; }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  if (which == GP_PROBE_COPY) {
    uint32_t ok = 0;
    int rc = copy_probe_locked(ctx, L.d, bytes, warmup, iters, t, checksum, &ok, nullptr);
    if (rc) return rc;
    if (!ok) { set_error("copy destination does not verify"); if (checksum) *checksum = ~*checksum; }
    return GP_OK;
  }
  return triad_probe_locked(ctx, L.d, bytes, warmup, iters, t, checksum, nullptr);
}

int gp_stream_host(gp_ctx* ctx, const char* uuid, uint32_t which, const void* b, const void* c,
                   float scalar, void* dst, uint64_t bytes) {
  if (which != GP_PROBE_COPY && which != GP_PROBE_TRIAD) { set_error("which must be COPY or TRIAD"); return GP_ERR_INVALID; }
  if ((bytes % 32) || (bytes && (!b || !dst)) || (which == GP_PROBE_TRIAD && bytes && !c)) {
    set_error("bytes must be a multiple of 32 and pointers non-null");
    return GP_ERR_INVALID;
  }
  if (bytes == 0) return GP_OK;
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  Device* d = L.d;
  const uint64_t pitch = (bytes + (1u << 20) - 1) & ~(uint64_t)((1u << 20) - 1);
  int rc = ensure_arena(ctx, d, 3 * pitch);
  if (rc) return rc;
  d->triad_operands_bytes = 0;
  const int variant = resolve_variant(ctx);
  uint8_t* A = d->arena;
  uint8_t* B = d->arena + pitch;
  uint8_t* C = d->arena + 2 * pitch;
  CU_OR_RETURN(cudaMemsetAsync(A, 0xEE, bytes, d->stream), nullptr);
  CU_OR_RETURN(cudaMemcpyAsync(B, b, bytes, cudaMemcpyHostToDevice, d->stream), nullptr);
  if (which == GP_PROBE_TRIAD) {
    CU_OR_RETURN(cudaMemcpyAsync(C, c, bytes, cudaMemcpyHostToDevice, d->stream), nullptr);
    CU_OR_RETURN(launch_triad(reinterpret_cast<float*>(A), reinterpret_cast<const float*>(B), reinterpret_cast<const float*>(C), scalar, bytes, variant, d->sm_count, d->stream), nullptr);
  } else {
    CU_OR_RETURN(launch_copy(A, B, bytes, variant, d->sm_count, d->stream), nullptr);
  }
  CU_OR_RETURN(cudaMemcpyAsync(dst, A, bytes, cudaMemcpyDeviceToHost, d->stream), nullptr);
  CU_OR_RETURN(cudaStreamSynchronize(d->stream), nullptr);
  return GP_OK;
}

int gp_gemm_probe(gp_ctx* ctx, const char* uuid, uint32_t mnk, uint64_t seed, uint32_t warmup,
                  uint32_t iters, uint32_t samples, gp_gemm_result* out) {
  if (!out || iters == 0) { set_error("null out or iters == 0"); return GP_ERR_INVALID; }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  return gemm_probe_locked(ctx, L.d, mnk ? mnk : 8192, seed ? seed : 0x5EEDull, warmup, iters, samples, out);
}

int gp_gemm_host(gp_ctx* ctx, const char* uuid, const uint16_t* a, const uint16_t* b, float* c,
                 uint32_t m, uint32_t n, uint32_t k) {
  if (!a || !b || !c) { set_error("null pointer"); return GP_ERR_INVALID; }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  return gemm_host_locked(ctx, L.d, a, b, c, m, n, k);
}

int gp_hbm_step(gp_ctx* ctx, const char* uuid, uint64_t bytes, uint32_t step, gp_pattern_result* out) {
  if (bytes == 0 || (bytes & ((1u << 20) - 1))) { set_error("bytes must be a non-zero multiple of 1 MiB"); return GP_ERR_INVALID; }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  Device* d = L.d;
  int rc = ensure_arena(ctx, d, 5 * bytes);
  if (rc) return rc;
  const int variant = resolve_variant(ctx);
  const uint64_t seed = ctx->cfg.seed ? ctx->cfg.seed : 0x5EEDull;
  const uint64_t nwords = bytes / 8, nf = bytes / 4;
  uint64_t* X = reinterpret_cast<uint64_t*>(d->arena);
  uint64_t* Y = reinterpret_cast<uint64_t*>(d->arena + bytes);
  float* Z = reinterpret_cast<float*>(d->arena + 2 * bytes);
  float* Bv = reinterpret_cast<float*>(d->arena + 3 * bytes);
  float* Cv = reinterpret_cast<float*>(d->arena + 4 * bytes);
  cudaStream_t s = d->stream;
  if (d->triad_operands_bytes != bytes || d->triad_operands_seed != seed) {
    CU_OR_RETURN(cudaEventRecord(d->events[0], s), nullptr);
    CU_OR_RETURN(launch_gen_f32(Bv, nf, seed ^ kSeedTriadBXor, d->sm_count, s), nullptr);
    CU_OR_RETURN(launch_gen_f32(Cv, nf, seed ^ kSeedTriadCXor, d->sm_count, s), nullptr);
    CU_OR_RETURN(cudaEventRecord(d->events[1], s), nullptr);
    CU_OR_RETURN(cudaStreamSynchronize(s), nullptr);
    float ms = 0;
    CU_OR_RETURN(cudaEventElapsedTime(&ms, d->events[0], d->events[1]), nullptr);
    d->stats[GP_K_INIT].launches += 2;
    d->stats[GP_K_INIT].total_ms += ms;
    d->stats[GP_K_INIT].algo_bytes += 2.0 * bytes;
    d->triad_operands_bytes = bytes;
    d->triad_operands_seed = seed;
  }
  const int wo_pass = (int)(step & 63), ad_pass = (int)(step & 1);
  cudaEvent_t* ev = d->events;
  CU_OR_RETURN(cudaEventRecord(ev[0], s), nullptr);
  CU_OR_RETURN(launch_pattern_fill(X, nwords, GP_PAT_WALKING_ONES, wo_pass, d->sm_count, s), nullptr);
  CU_OR_RETURN(cudaEventRecord(ev[1], s), nullptr);
  CU_OR_RETURN(launch_copy(Y, X, bytes, variant, d->sm_count, s), nullptr);
  CU_OR_RETURN(cudaEventRecord(ev[2], s), nullptr);
  CU_OR_RETURN(launch_pattern_verify(Y, nwords, GP_PAT_WALKING_ONES, wo_pass, d->scratch, d->d_results + 0, d->sm_count, s), nullptr);
  CU_OR_RETURN(cudaEventRecord(ev[3], s), nullptr);
  CU_OR_RETURN(launch_pattern_fill(X, nwords, GP_PAT_ADDRESS, ad_pass, d->sm_count, s), nullptr);
  CU_OR_RETURN(cudaEventRecord(ev[4], s), nullptr);
  CU_OR_RETURN(launch_pattern_verify(X, nwords, GP_PAT_ADDRESS, ad_pass, d->scratch, d->d_results + 1, d->sm_count, s), nullptr);
  CU_OR_RETURN(cudaEventRecord(ev[5], s), nullptr);
  CU_OR_RETURN(launch_triad(Z, Bv, Cv, 3.0f, bytes, variant, d->sm_count, s), nullptr);
  CU_OR_RETURN(cudaEventRecord(ev[6], s), nullptr);
  CU_OR_RETURN(launch_pattern_verify(reinterpret_cast<uint64_t*>(Z), nwords, kPatNone, 0, d->scratch, d->d_results + 2, d->sm_count, s), nullptr);
  CU_OR_RETURN(cudaEventRecord(ev[7], s), nullptr);
  CU_OR_RETURN(cudaMemcpyAsync(d->h_results, d->d_results, 3 * sizeof(gp_pattern_result), cudaMemcpyDeviceToHost, s), nullptr);
  CU_OR_RETURN(cudaStreamSynchronize(s), nullptr);
  static const int kid[7] = {GP_K_FILL, GP_K_COPY, GP_K_VERIFY, GP_K_FILL, GP_K_VERIFY, GP_K_TRIAD, GP_K_VERIFY};
  static const double mult[7] = {1, 2, 1, 1, 1, 3, 1};
  for (int i = 0; i < 7; ++i) {
    float ms = 0;
    CU_OR_RETURN(cudaEventElapsedTime(&ms, ev[i], ev[i + 1]), nullptr);
    d->stats[kid[i]].launches += 1;
    d->stats[kid[i]].total_ms += ms;
    d->stats[kid[i]].algo_bytes += mult[i] * (double)bytes;
  }
  {
    float ms = 0;
    CU_OR_RETURN(cudaEventElapsedTime(&ms, ev[0], ev[7]), nullptr);
    d->stats[GP_K_STEP].launches += 1;
    d->stats[GP_K_STEP].total_ms += ms;
    d->stats[GP_K_STEP].algo_bytes += 10.0 * (double)bytes;
  }
  if (out) memcpy(out, d->h_results, 3 * sizeof(gp_pattern_result));
  return GP_OK;
}

int gp_kernel_stats(gp_ctx* ctx, const char* uuid, gp_kernel_stat* out, int reset) {
  Device* d = find_device(ctx, uuid);
  if (!d) return GP_ERR_NO_DEVICE;
  std::lock_guard<std::mutex> g(d->mu);
  if (out) memcpy(out, d->stats, sizeof(d->stats));
  if (reset) memset(d->stats, 0, sizeof(d->stats));
  return GP_OK;
}

int gp_inject_bitflip(gp_ctx* ctx, const char* uuid, uint64_t word_index, int bit) {
  if (bit < 0 || bit > 63) { set_error("bit out of range"); return GP_ERR_INVALID; }
  Locked L(ctx, uuid);
  if (L.rc) return L.rc;
  if ((word_index + 1) * 8 > L.d->arena_bytes) { set_error("word index outside the arena"); return GP_ERR_INVALID; }
  CU_OR_RETURN(launch_bitflip(reinterpret_cast<uint64_t*>(L.d->arena), word_index, bit, L.d->stream), nullptr);
  CU_OR_RETURN(cudaStreamSynchronize(L.d->stream), nullptr);
  return GP_OK;
}

int gp_probe_p2p(gp_ctx* ctx, int ngpus, size_t bytes, int dir, double* gbps_out) {
  if (!ctx || !gbps_out || ngpus < 1 || dir < 0 || dir > 6) { set_error("bad arguments"); return GP_ERR_INVALID; }
  return p2p_probe(ctx, ngpus, bytes, dir, gbps_out);
}

int gp_probe_fabric(gp_ctx* ctx, int ngpus, gp_fabric_result* out) {
  if (!ctx || !out || ngpus < 1 || ngpus > GP_FABRIC_MAX_GPUS) { set_error("bad arguments (1 <= ngpus <= %d)", GP_FABRIC_MAX_GPUS); return GP_ERR_INVALID; }
  return fabric_probe(ctx, ngpus, out);
}

int gp_probe_allreduce(gp_ctx* ctx, int nranks, size_t bytes, uint32_t warmup, uint32_t iters,
                       double* algbw_gbps, double* busbw_gbps) {
  if (!ctx || nranks < 1 || iters == 0) { set_error("bad arguments"); return GP_ERR_INVALID; }
  return allreduce_probe(ctx, nranks, bytes, warmup, iters, algbw_gbps, busbw_gbps);
}

}  // extern "C"
