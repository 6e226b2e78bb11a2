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
 *   - every entry point is blocking (one exception: with keep_arena = 0 a probe
 *     returns while its device memory is still being released on a background
 *     thread; anything that needs the allocator on that device waits for it),
 *     takes the per-device lock, and calls cudaSetDevice on entry: safe from arbitrary OS threads (cgo calls land
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

#define GP_ABI_VERSION 2u

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
  GP_ERR_TOO_LARGE = -8,    /* request exceeds a fixed compile-time limit */
  GP_ERR_DEVICE_LOST = -9   /* a sticky CUDA fault (illegal address, uncorrectable ECC, Xid ...) has killed this
                             * process's context on the device: every later call on it returns this code until
                             * gp_device_reset, or until the device is graded again from a fresh process */
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

/* GEMM kernels (gp_config.gemm_variant) */
enum {
  GP_GEMM_AUTO = 0,           /* 2-CTA pairs; 256 x 512 tiles when N % 512 == 0, else 256 x 256 */
  GP_GEMM_1CTA_128x256 = 1,   /* cta_group::1, one CTA per 128 x 256 tile */
  GP_GEMM_2CTA_256x256 = 2,   /* cta_group::2, 256 x 256 tile per pair, accumulator double-buffered */
  GP_GEMM_2CTA_256x512 = 3    /* as AUTO */
};
/* test-only fault injection (gp_config.test_fault); nothing here touches the hardware abnormally */
enum {
  GP_TEST_FAULT_GEMM_SKIP_TILE = 1u << 0, /* after the launches, one 256 x 256 tile of C is put back to its NaN fill */
  GP_TEST_FAULT_GEMM_FLIP_BIT = 1u << 1,  /* one low mantissa bit of one element of the timed C is flipped */
  GP_TEST_FAULT_P2P_CORRUPT = 1u << 2     /* one bit of one word of the last GPU's P2P source buffer is flipped after the
                                           * fill: whoever reads it must notice (gp_probe_p2p fails, the fabric verdict
                                           * reports data_ok = 0) */
};

/* arena backends (gp_config.arena_backend) */
enum {
  GP_ARENA_AUTO = 0,   /* keep_arena = 1: VMM when the driver offers it; keep_arena = 0: cudaMalloc (cheapest cold) */
  GP_ARENA_VMM = 1,    /* one reserved virtual range per device, 4 GiB physical chunks (cuMemCreate) mapped as the
                        * arena grows: growing never moves or re-allocates what is already mapped */
  GP_ARENA_MALLOC = 2  /* cudaMalloc: one allocation per size change for a kept arena; for a cold gp_probe_device
                        * (keep_arena = 0) three slabs allocated in order behind the probes that do not need them yet */
};
#define GP_FABRIC_MAX_GPUS 16

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
  uint32_t keep_arena;     /* 1: keep device memory between probes (gp_trim / gp_trim_device give it back);
                            * 0: gp_probe_* hand it to a release thread as they return */
  uint32_t gemm_output;    /* GEMM probe's C: 0 = fp32 (strictest check), 1 = bf16 (what cuBLAS bf16 peaks are quoted on) */
  uint64_t seed;           /* GEMM / triad operand generator seed; 0 -> 0x5EED */
  double min_copy_gbps;    /* health floors; 0 -> not enforced */
  double min_triad_gbps;
  double min_gemm_tflops;
  double gemm_tol;         /* |c-ref| <= gemm_tol * sum_k|a||b| against the fp64 host reference; 0 -> 1e-3.
                            * A bf16 C is compared, bit for bit, with RNE(bf16) of the fp32 C of the same problem. */
  /* ---- ABI 2 ---- */
  double min_p2p_gbps;     /* fabric floors for gp_probe_fabric; 0 -> not enforced.  Worst off-diagonal pair, GB/s */
  double max_p2p_spread;   /* largest allowed max/min over the off-diagonal pairs (a flat matrix is the NVSwitch
                            * health signal); 0 -> not enforced */
  double min_allreduce_busbw_gbps;
  uint64_t fabric_bytes;   /* bytes per transfer of gp_probe_fabric; 0 -> 1 GiB */
  uint32_t arena_backend;  /* GP_ARENA_* */
  uint32_t sm_limit;       /* size every persistent grid as if the device had this many SMs (a MIG slice exposes
                            * 1/7 .. 4/7 of them; tests use this to run the launchers at slice scale); 0 -> all */
  uint32_t gemm_variant;   /* GP_GEMM_*: which tcgen05 kernel the GEMM probe launches */
  uint32_t test_fault;     /* GP_TEST_FAULT_* bits, tests only: make a check see what a broken GPU would produce */
  uint32_t reserved2[2];
  uint64_t triad_expect_sum;  /* known answer for the triad probe at this hbm_bytes / seed: SUM and XOR (mod 2^64) of the */
  uint64_t triad_expect_xor;  /* u64 words of a = fmaf(3, c, b), computed OFF the device (the CPU oracle's
                               * gp_cpu_triad_checksum; INTEGRATION.md lists the default configuration's).  Both 0 ->
                               * not compared.  triad_ok alone is the GPU checking itself with the same FMA units. */
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
  uint32_t output_bf16;    /* 1 if the timed C was bf16.  The fp64 checks then run on the fp32 C of the same launch
                            * plan (one extra untimed launch) and the bf16 C must equal RNE(bf16) of it bit for bit */
  uint32_t reserved;
  uint64_t bf16_mismatches;/* elements of the bf16 C that differ from RNE(fp32 C); counted into samples_bad */
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
  /* ---- ABI 2: where duration_s went (host steady_clock) ---- */
  double alloc_s;           /* host time this call spent WAITING for memory: growing the arena, minus whatever part of
                             * that ran on the helper thread behind the GEMM probe (0 when the kept arena was large enough) */
  double free_s;            /* releasing it before return; 0 when the release was handed to the reaper thread */
  double kernels_s;         /* duration_s - alloc_s - free_s: generation, warm-up, timed launches, checks, copies */
  uint64_t arena_bytes;     /* device memory the call ran in */
  uint32_t arena_reused;    /* 1 if the arena was already there on entry (periodic re-probe of an idle device) */
  uint32_t arena_backend;   /* GP_ARENA_VMM or GP_ARENA_MALLOC actually used */
  uint64_t triad_sum, triad_xor; /* SUM / XOR of the u64 words of the triad's result array */
  uint32_t triad_known_answer;   /* 0 not compared (no expectation configured), 1 equal, 2 different (=> TRIAD failed) */
  uint32_t free_deferred;   /* 1 if keep_arena = 0 and the arena is being released on a background thread as the call
                             * returns (the memory is back with the driver a moment later; the next call waits for it) */
} gp_device_result;

/* Verdict on the GPU-to-GPU fabric (K4 of SURVEY.md §8): the P2P matrices, their spread, the worst pair and the
 * NCCL all-reduce bus bandwidth, against the floors of gp_config.  Row = source GPU, n*n entries used. */
typedef struct gp_fabric_result {
  uint32_t struct_size;
  uint32_t ngpus;
  uint64_t bytes;                /* per transfer */
  double read_gbps[GP_FABRIC_MAX_GPUS * GP_FABRIC_MAX_GPUS];   /* kernel on row GPU loads from column GPU */
  double write_gbps[GP_FABRIC_MAX_GPUS * GP_FABRIC_MAX_GPUS];  /* kernel on row GPU stores to column GPU  */
  double read_min, read_max, read_mean;     /* off-diagonal */
  double write_min, write_max, write_mean;
  double spread;                 /* max over {read, write} of max/min off-diagonal; 1.0 = perfectly flat */
  int32_t worst_src, worst_dst;  /* the slowest ordered pair over both matrices, -1 if ngpus < 2 */
  double worst_gbps;
  double allreduce_algbw_gbps, allreduce_busbw_gbps;  /* 0 if the all-reduce was not run (ngpus < 2 or no NCCL) */
  int32_t allreduce_error;       /* gp_error of the all-reduce leg (NCCL missing is reported here, not fatal) */
  uint32_t data_ok;              /* 1 iff every peer read returned the bytes the peer holds and every peer write landed */
  uint32_t failed;               /* bit 0: below min_p2p_gbps, 1: spread above max_p2p_spread, 2: all-reduce below its
                                  * floor, 3: data mismatch */
  int32_t healthy;               /* 1 iff failed == 0 */
  double duration_s;
} gp_fabric_result;

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
/* Free ONE device's arena.  With keep_arena = 1 the arena of an idle device stays mapped between periodic
 * re-probes (no allocation inside the probe call); the plugin hands the memory back the moment kubelet
 * assigns the device: call this from Allocate (plugin/plugin.go:210-225), before the pod starts. */
int gp_trim_device(gp_ctx* ctx, const char* uuid);
/* Recover a device that returned GP_ERR_DEVICE_LOST without restarting the process: drops every handle the
 * library holds on it and calls cudaDeviceReset.  This destroys the device's primary context for the WHOLE
 * process, so it is only for hosts where nothing else uses CUDA on that device (the Go plugin is one).  The
 * reference's own recovery is a full plugin restart (plugin/plugin.go:107-129). */
int gp_device_reset(gp_ctx* ctx, const char* uuid);
/* 1 if a sticky fault has been seen on the device since the last reset, else 0; negative gp_error on bad arguments. */
int gp_device_poisoned(gp_ctx* ctx, const char* uuid);
uint32_t gp_abi_version(void);
/* sizeof() of the POD structs as this library was compiled, so a binding in
 * another language can check its own layout: which = 0 gp_config,
 * 1 gp_pattern_result, 2 gp_timing, 3 gp_gemm_result, 4 gp_device_result,
 * 5 gp_kernel_stat, 6 gp_nvml_info, 7 gp_fabric_result; 0 for anything else. */
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
 *   - d_words must be 32-byte aligned (the kernels load 256 bits at a time) and d_result
 *     8-byte aligned, else GP_ERR_INVALID: grading a sub-range of a buffer is fine as long as
 *     it starts on a 32-byte boundary;
 *   - the address looked up on the first call is cached per device ordinal for the life of the
 *     process.  cudaDeviceReset destroys the array with the rest of the context: gp_device_reset
 *     drops the cached address, but after a reset done behind the library's back this entry
 *     point must not be called again on that device;
 *   - calls cudaSetDevice(device_ordinal) on the calling thread. */
#define GP_NOALLOC_MAX_WORDS (1ull << 35) /* 256 GiB of graded memory per call */
int gp_pattern_verify_noalloc(int device_ordinal, const uint64_t* d_words, uint64_t nwords, int kind,
                              int pass, gp_pattern_result* d_result, void* stream);

/* Test-only fault injection: flip `bit` of arena word `word_index`. */
int gp_inject_bitflip(gp_ctx* ctx, const char* uuid, uint64_t word_index, int bit);
/* Test-only: behave as if a CUDA call on the device had just failed with `cuda_error` (a cudaError_t value).
 * Nothing is launched and the GPU is not touched; the error goes through the library's own classification, so
 * a sticky code (e.g. 700 cudaErrorIllegalAddress, 214 cudaErrorECCUncorrectable) marks the device lost exactly
 * as the real fault would.  Returns the gp_error the failing call would have returned. */
int gp_inject_cuda_error(gp_ctx* ctx, const char* uuid, int cuda_error);

/* ---- multi-GPU (the only cross-device paths) ---------------------------- */

/* dir: 0 = read (kernel on src loads from dst's memory), 1 = write (kernel on
 * src stores to dst), 2 = bidirectional write (src->dst and dst->src at once).
 * gbps is n*n row-major, row = src; diagonal = local bandwidth.
 * dir 3 = all pairs at once (switch saturation): every GPU reads `bytes` from
 * each peer concurrently; row i = GPU i's aggregate ingress / (n-1), diagonal =
 * the aggregate.  dir 4 = bidirectional read (both GPUs of a pair read from each
 * other at once).  dir 5 / 6 = the copy-engine yardstick: cudaMemcpyPeerAsync
 * src -> dst, one direction / both directions at once, same schedule and timing
 * as the kernels so the two can be read side by side.
 * Every read mode folds what it loaded (sum and xor of the u64 words) and the
 * host compares that with the closed form of the address pattern the peer
 * holds; every write mode checks what landed.  A mismatch is GP_ERR_CUDA. */
enum {
  GP_P2P_READ = 0, GP_P2P_WRITE = 1, GP_P2P_BIDIR_WRITE = 2, GP_P2P_ALL_AT_ONCE = 3,
  GP_P2P_BIDIR_READ = 4, GP_P2P_COPY_ENGINE = 5, GP_P2P_COPY_ENGINE_BIDIR = 6
};
int gp_probe_p2p(gp_ctx* ctx, int ngpus, size_t bytes, int dir, double* gbps);
/* Read + write matrices at gp_config.fabric_bytes, then the all-reduce, graded against the fabric floors.
 * Called once after the single-GPU grade (plugin/manager.go:156-174); its matrix is what
 * GetPreferredAllocation ranks by (plugin/plugin.go:192-207, :256-282). */
int gp_probe_fabric(gp_ctx* ctx, int ngpus, gp_fabric_result* out);
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
