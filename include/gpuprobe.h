/*
 * gpuprobe.h — C ABI of libgpuprobe.so, the B200-native per-device health and
 * benchmark probe that sits behind uppercaveman/k8s-gpu-device-plugin.
 *
 * The reference has NO FFI boundary of its own on this path (its only cgo is
 * inside third-party go-nvml, reference go.mod:8) and NO probe arithmetic
 * (benchmark/benchmark.go:14-124 is a runtime/pprof wrapper;
 * device/devices.go:74 hard-codes Health = Healthy).  Each entry point below
 * therefore cites the reference call site that a maintainer would route
 * through it, not a function it re-implements.  The frozen spec for the
 * arithmetic is SURVEY.md §8(c); the CPU restatement lives in oracle/ and is
 * test infrastructure only — nothing in this library calls it.
 *
 * Conventions
 *   - return 0 (GP_OK) on success, a negative gp_error otherwise; the raw
 *     CUDA / NCCL code is kept in the result struct's `detail` field and in
 *     gp_last_error_detail().
 *   - caller owns every output buffer; the library never retains a caller
 *     pointer past the call (cgo rule: no Go pointers held by C).
 *   - the library owns gp_ctx and all device memory.
 *   - every entry point is blocking, takes the per-device lock, and calls
 *     cudaSetDevice on entry: safe from arbitrary OS threads (cgo calls land
 *     on arbitrary threads).  Calls for DIFFERENT devices run concurrently;
 *     calls for the same device serialise.
 *   - no stdout/stderr, no signals, no exit().
 *   - result structs are POD with fixed-width fields; the first field is
 *     struct_size so the Go side can detect skew.
 *   - devices are addressed by NVML-style UUID string ("GPU-xxxxxxxx-xxxx-…",
 *     the key reference device/devices.go:71 uses); the decimal CUDA ordinal
 *     ("0", "1", …) is accepted as well for harnesses.
 */
#ifndef GPUPROBE_H_
#define GPUPROBE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
/* the library is built with -fvisibility=hidden; only this header's symbols are exported */
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define GP_ABI_VERSION 1u

typedef struct gp_ctx gp_ctx;

typedef enum gp_error {
  GP_OK = 0,
  GP_ERR_INVALID = -1,      /* bad argument (NULL, size not a multiple of the unit, unknown kind) */
  GP_ERR_NO_DEVICE = -2,    /* uuid / ordinal not visible to CUDA */
  GP_ERR_CUDA = -3,         /* CUDA runtime/driver error, code in detail */
  GP_ERR_NCCL = -4,         /* NCCL error or libnccl not loadable, code in detail */
  GP_ERR_NOMEM = -5,        /* arena allocation failed / request exceeds arena cap */
  GP_ERR_UNSUPPORTED = -6,  /* device is not sm_100 (tcgen05/TMA paths) or no peer access */
  GP_ERR_STRUCT_SIZE = -7,  /* struct_size mismatch between caller and library */
  GP_ERR_TOO_LARGE = -8     /* request exceeds a fixed compile-time limit */
} gp_error;

/* pattern kinds — SURVEY.md §8(c) */
enum {
  GP_PAT_WALKING_ONES = 0,  /* pass p in [0,64):  P(i) = 1 << ((i + p) & 63)      */
  GP_PAT_WALKING_ZEROS = 1, /* pass p in [0,64):  P(i) = ~(1 << ((i + p) & 63))   */
  GP_PAT_ADDRESS = 2        /* pass 0: P(i) = 8*i ; pass 1: P(i) = ~(8*i)         */
};

/* probe selection mask for gp_probe_device / gp_probe_all */
enum {
  GP_PROBE_COPY = 1u << 0,
  GP_PROBE_TRIAD = 1u << 1,
  GP_PROBE_PATTERN = 1u << 2,
  GP_PROBE_GEMM = 1u << 3,
  GP_PROBE_ALL_SINGLE = 0xFu
};

/* kernel ids for gp_kernel_stats */
enum {
  GP_K_COPY = 0,
  GP_K_TRIAD = 1,
  GP_K_FILL = 2,
  GP_K_VERIFY = 3,
  GP_K_GEMM = 4,
  GP_K_P2P = 5,
  GP_K_INIT = 6, /* input generators (triad operands, GEMM operands) */
  GP_K_STEP = 7, /* one whole gp_hbm_step, first event to last (not a kernel) */
  GP_K_COUNT = 8
};

/* stream kernel variants (gp_config.stream_variant) */
enum {
  GP_STREAM_AUTO = 0,
  GP_STREAM_VEC = 1, /* 256-bit ld/st per thread, register staged            */
  GP_STREAM_TMA = 2  /* cp.async.bulk global->smem->global, mbarrier pipeline */
};

#define GP_PATTERN_ALL_FREE UINT64_MAX

typedef struct gp_config {
  uint32_t struct_size;    /* = sizeof(gp_config) */
  uint32_t abi_version;    /* = GP_ABI_VERSION */
  uint64_t hbm_bytes;      /* bytes per stream array; 0 -> 16 GiB (BASELINE.json configs[1]) */
  uint64_t pattern_bytes;  /* bytes graded by fill/verify; 0 -> hbm_bytes; GP_PATTERN_ALL_FREE -> every free
                            * byte of the device but a 1 GiB margin (whole-memory grading) */
  uint64_t arena_cap_bytes;/* hard cap for the per-device arena; 0 -> 5*hbm_bytes + 1 GiB */
  uint32_t gemm_mnk;       /* M=N=K of the GEMM probe; 0 -> 8192 */
  uint32_t warmup;         /* untimed launches per kernel; 0 -> 2 */
  uint32_t iters;          /* timed launches per kernel; 0 -> 10 */
  uint32_t wo_passes;      /* walking-ones passes [0,64]; 0 -> 64 */
  uint32_t gemm_samples;   /* sampled C entries checked on the host in fp64; 0 -> 4096 */
  uint32_t stream_variant; /* GP_STREAM_* */
  uint32_t keep_arena;     /* 1: keep device memory between probes; 0: free after gp_probe_* */
  uint32_t gemm_output;    /* GEMM probe's C: 0 = fp32 (strictest check), 1 = bf16 (what cuBLAS bf16 peaks are quoted on) */
  uint64_t seed;           /* GEMM / triad operand generator seed; 0 -> 0x5EED */
  double min_copy_gbps;    /* health floors; 0 -> not enforced */
  double min_triad_gbps;
  double min_gemm_tflops;
  double gemm_tol;         /* |c-ref| <= gemm_tol * sum_k|a||b| (+ 2^-8 |ref| for bf16 output); 0 -> 1e-3 */
} gp_config;

/* One verify pass — SURVEY.md §8(c).  All sums are mod 2^64. */
typedef struct gp_pattern_result {
  uint64_t sum;             /* SUM w[i]                      */
  uint64_t xor_;            /* XOR w[i]                      */
  uint64_t wsum;            /* SUM w[i] * (2i + 1)           */
  uint64_t mismatches;      /* #{i : w[i] != P(i)}           */
  uint64_t first_bad_index; /* min such i, UINT64_MAX if none */
  uint64_t first_bad_xor;   /* w[i] ^ P(i) at that i, 0 if none */
} gp_pattern_result;

typedef struct gp_timing {
  double best_ms;
  double median_ms;
  double mean_ms;
  uint32_t launches;
  uint32_t reserved;
} gp_timing;

typedef struct gp_gemm_result {
  uint32_t struct_size;
  uint32_t m, n, k;
  gp_timing t;
  double tflops_best;
  double tflops_median;
  double max_scaled_err;   /* max over samples of |c-ref| / sum_k|a_ik||b_jk| */
  double freivalds_err;    /* max_i |(C x)_i - (A (B^T x))_i| / (sum_j sum_k |a||b||x|) */
  uint32_t samples;
  uint32_t samples_bad;    /* samples exceeding gemm_tol */
  uint64_t c_checksum;     /* XOR of C's bit patterns folded to 64 bit */
  uint32_t output_bf16;    /* 1 if C was bf16: errors are then measured beyond the 2^-8 |ref| output rounding */
  uint32_t reserved;
} gp_gemm_result;

typedef struct gp_device_result {
  uint32_t struct_size;     /* = sizeof(gp_device_result) */
  uint32_t abi_version;
  char uuid[48];            /* "GPU-…" NUL-terminated */
  int32_t ordinal;          /* CUDA ordinal */
  int32_t sm_count;
  uint64_t total_mem;       /* bytes */
  int32_t cc_major, cc_minor;
  uint32_t probed_mask;     /* GP_PROBE_* actually run */
  uint32_t failed_mask;     /* GP_PROBE_* whose verdict failed */
  int32_t healthy;          /* 1 iff failed_mask == 0 and error == 0 */
  int32_t error;            /* gp_error of the first failing call */
  int64_t detail;           /* raw cudaError_t / ncclResult_t */
  double duration_s;        /* host steady_clock, entry to return */
  uint64_t hbm_bytes;       /* bytes per stream array used */
  uint64_t pattern_bytes;
  gp_timing copy_t;  double copy_gbps_best, copy_gbps_median;
  gp_timing triad_t; double triad_gbps_best, triad_gbps_median;
  gp_timing fill_t;  double fill_gbps_best, fill_gbps_median;
  gp_timing verify_t;double verify_gbps_best, verify_gbps_median;
  uint64_t copy_checksum;   /* sum of dst words; equals the source sum */
  uint32_t copy_ok, triad_ok;
  uint32_t pattern_passes;  /* fill+verify pairs run */
  uint32_t pattern_bad_passes;
  uint64_t pattern_mismatches;
  int32_t first_bad_kind, first_bad_pass;
  uint64_t first_bad_index, first_bad_xor;
  uint64_t pattern_digest;  /* checksum of checksums: SUM over passes of (sum ^ rotl(xor,1) ^ rotl(wsum,2)) * (pass_no+1) */
  gp_gemm_result gemm;
} gp_device_result;

typedef struct gp_kernel_stat {
  uint64_t launches;
  double total_ms;          /* CUDA-event time on the library stream */
  double algo_bytes;        /* algorithmic bytes (or flops for GP_K_GEMM) summed over launches */
} gp_kernel_stat;

/* ---- lifecycle ---------------------------------------------------------- */

/* Replaces nothing; new.  Called once where the reference builds its NVML
 * handle (plugin/manager.go:44).  cfg may be NULL for defaults. */
int gp_init(const gp_config* cfg, gp_ctx** out);
void gp_shutdown(gp_ctx* ctx);
/* Free every per-device arena (pods need the memory once grading is done). */
int gp_trim(gp_ctx* ctx);
uint32_t gp_abi_version(void);
/* sizeof() of the POD structs as this library was compiled, so a binding in
 * another language can check its own layout: which = 0 gp_config,
 * 1 gp_pattern_result, 2 gp_timing, 3 gp_gemm_result, 4 gp_device_result,
 * 5 gp_kernel_stat, 6 gp_nvml_info; 0 for anything else. */
size_t gp_sizeof(int which);
const char* gp_strerror(int err);
/* Thread-local text of the last failure on this thread ("" if none). */
const char* gp_last_error_detail(void);

/* ---- enumeration (mirrors device/device.go:37-43 GetUUID) ---------------- */
int gp_device_count(gp_ctx* ctx);
int gp_device_uuid(gp_ctx* ctx, int ordinal, char* buf, size_t buflen);

/* Clocks, power, throttle reasons and — what the plugin gates its periodic re-probe on — how
 * many OTHER processes are computing on the device right now.  Read through NVML
 * (dlopen libnvidia-ml.so.1, the library go-nvml binds, reference go.mod:8); when it is not
 * loadable the call still returns GP_OK with available = 0. */
typedef struct gp_nvml_info {
  uint32_t struct_size;
  uint32_t available;          /* 1 iff NVML answered for this device */
  uint32_t sm_clock_mhz, sm_clock_max_mhz, mem_clock_mhz;
  uint32_t power_mw, temperature_c;
  uint32_t other_compute_procs;/* compute processes on the device, not counting the caller */
  uint64_t throttle_reasons;   /* nvmlClocksEventReasons bit mask */
  uint64_t mem_used_bytes, mem_total_bytes;
} gp_nvml_info;
int gp_nvml_snapshot(gp_ctx* ctx, const char* uuid, gp_nvml_info* out);

/* ---- the probe: verdict that replaces `dev.Health = pluginapi.Healthy`
 *      (device/devices.go:74); dispatched from the device-map build loop
 *      (device/device_map.go:48-75, :101-111) ------------------------------- */
int gp_probe_device(gp_ctx* ctx, const char* uuid, uint32_t which_mask, gp_device_result* out);
/* All single-device probes concurrently on the first ngpus devices, one host
 * thread per GPU (plugin/manager.go:156-174 loadPlugins is the caller).
 * out has ngpus entries; wall_s is entry-to-return steady_clock seconds. */
int gp_probe_all(gp_ctx* ctx, int ngpus, uint32_t which_mask, gp_device_result* out, double* wall_s);

/* ---- building blocks (parity tests and bench.py drive these) ------------- */

/* Fill arena words [0, nwords) with pattern (kind, pass).  nwords % 64 == 0. */
int gp_pattern_fill(gp_ctx* ctx, const char* uuid, int kind, int pass, uint64_t nwords);
/* Verify arena words [0, nwords) against (kind, pass). */
int gp_pattern_verify(gp_ctx* ctx, const char* uuid, int kind, int pass, uint64_t nwords,
                      gp_pattern_result* out);
/* Host-buffer form, the GPU twin of the oracle's gp_cpu_pattern: copies
 * host_words (may be pageable) to the arena, verifies, returns the result.
 * If fill != 0 the device fills first and the words are copied back to
 * host_words after the verify, as the CPU checker leaves them. */
int gp_pattern_host(gp_ctx* ctx, const char* uuid, uint64_t* host_words, uint64_t nwords,
                    int kind, int pass, int fill, gp_pattern_result* out);
/* passes [pass_begin, pass_begin+pass_count) of `kind`, fill+verify each,
 * results[pass_count]; timings are per launch. */
int gp_pattern_probe(gp_ctx* ctx, const char* uuid, int kind, int pass_begin, int pass_count,
                     uint64_t nwords, gp_pattern_result* results, gp_timing* fill_t,
                     gp_timing* verify_t);

/* which = GP_PROBE_COPY or GP_PROBE_TRIAD; bytes per array, multiple of 1 MiB.
 * checksum: copy -> SUM of dst u64 words; triad -> count of fp32 elements
 * that differ from fmaf(s, c, b) recomputed on the device from the
 * generator (0 when healthy). */
int gp_stream_probe(gp_ctx* ctx, const char* uuid, uint32_t which, uint64_t bytes,
                    uint32_t warmup, uint32_t iters, gp_timing* t, uint64_t* checksum);
/* Run copy / triad once over caller data (parity tests): src arrays are
 * copied to the arena, the kernel runs once, dst is copied back.  bytes is a
 * multiple of 32; for copy, b is the source and c is ignored. */
int gp_stream_host(gp_ctx* ctx, const char* uuid, uint32_t which, const void* b, const void* c,
                   float scalar, void* dst, uint64_t bytes);

/* M = N = K = mnk (a multiple of 256; 0 -> 8192) from the counter generator; timed launches,
 * then `samples` stratified entries checked against an fp64 host reference plus a Freivalds
 * check of the whole product.  C is fp32 or bf16 per gp_config.gemm_output. */
int gp_gemm_probe(gp_ctx* ctx, const char* uuid, uint32_t mnk, uint64_t seed, uint32_t warmup,
                  uint32_t iters, uint32_t samples, gp_gemm_result* out);
/* C (fp32, row-major m x n) = A (bf16 bits, m x k, K-major) * B^T (bf16 bits,
 * n x k, K-major) on caller data (parity tests). m,n multiples of 256, k of 64. */
int gp_gemm_host(gp_ctx* ctx, const char* uuid, const uint16_t* a, const uint16_t* b, float* c,
                 uint32_t m, uint32_t n, uint32_t k);

/* One bench "step" over `bytes` per array (arena = 5 arrays X, Y, Z, b, c),
 * enqueued back-to-back on the device's stream, blocking until done:
 *   fill X walking-ones(pass = step & 63); copy X -> Y; verify Y  -> out[0]
 *   fill X address(pass = step & 1);       verify X                -> out[1]
 *   triad Z = b + 3 c (b, c generated once per size/seed);
 *   checksum of Z's u64 words (mismatches = 0)                      -> out[2]
 * Algorithmic bytes per step: 10 * bytes (fill 2, copy 2, verify 3, triad 3). */
int gp_hbm_step(gp_ctx* ctx, const char* uuid, uint64_t bytes, uint32_t step,
                gp_pattern_result* out /*[3]*/);
int gp_kernel_stats(gp_ctx* ctx, const char* uuid, gp_kernel_stat* out /*[GP_K_COUNT]*/, int reset);

/* Verify without allocating anything.  No counterpart in the reference.
 *
 * For callers that cannot call cudaMalloc: a re-grade of a GPU whose memory is
 * already leased to pods (an allocation there fails or eats into a tenant's
 * share), or code that must not enter the CUDA allocator because cudaMalloc and
 * cudaFree synchronise the device.  The caller passes the DEVICE memory to grade
 * (d_words, nwords u64 words, nwords % 64 == 0) and a DEVICE pointer for the
 * 48-byte result; the work is enqueued on `stream` (a cudaStream_t, NULL for the
 * default stream) of device `device_ordinal` and the call returns without
 * synchronising — wait on the stream before reading *d_result.  Same arithmetic
 * and same bits as gp_pattern_verify.
 *
 * The kernel's scratch (accumulator slots) is a fixed 33,024-byte array inside
 * the library's device image instead of allocated memory; every GPU that loads
 * the library holds it until the process exits.  Limits and rules:
 *   - at most GP_NOALLOC_MAX_WORDS words per call, else GP_ERR_TOO_LARGE;
 *   - the first call on each device looks the array's address up
 *     (cudaGetSymbolAddress); make that first call outside any stream capture;
 *   - all calls on a device share the one array: two calls whose kernels overlap
 *     on the GPU (different streams, or different threads) race and return
 *     garbage — serialise them; calls on one stream are safe;
 *   - calls cudaSetDevice(device_ordinal) on the calling thread. */
#define GP_NOALLOC_MAX_WORDS (1ull << 35) /* 256 GiB of graded memory per call */
int gp_pattern_verify_noalloc(int device_ordinal, const uint64_t* d_words, uint64_t nwords, int kind,
                              int pass, gp_pattern_result* d_result, void* stream);

/* Test-only fault injection: flip `bit` of arena word `word_index`. */
int gp_inject_bitflip(gp_ctx* ctx, const char* uuid, uint64_t word_index, int bit);

/* ---- multi-GPU (the only cross-device paths) ---------------------------- */

/* dir: 0 = read (kernel on src loads from dst's memory), 1 = write (kernel on
 * src stores to dst), 2 = bidirectional write (src->dst and dst->src at once).
 * gbps is n*n row-major, row = src; diagonal = local bandwidth.
 * dir 3 = all pairs at once (switch saturation): every GPU reads `bytes` from
 * each peer concurrently; row i = GPU i's aggregate ingress / (n-1), diagonal =
 * the aggregate.  dir 4 = bidirectional read (both GPUs of a pair read from each
 * other at once). */
int gp_probe_p2p(gp_ctx* ctx, int ngpus, size_t bytes, int dir, double* gbps);
/* In-place fp32 sum ncclAllReduce over the first nranks devices. */
int gp_probe_allreduce(gp_ctx* ctx, int nranks, size_t bytes, uint32_t warmup, uint32_t iters,
                       double* algbw_gbps, double* busbw_gbps);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* GPUPROBE_H_ */
