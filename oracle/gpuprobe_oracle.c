/*
 * gpuprobe_oracle.c — CPU restatement of the probe arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this file's
 * library.  Nothing under k8s-gpu-device-plugin_b200/ links, loads or calls
 * it; the product path fails loudly without its CUDA library instead.
 *
 * PARITY UNPINNED.  The reference (uppercaveman/k8s-gpu-device-plugin) holds
 * no implementation of this path and no tests at all:
 *   - benchmark/benchmark.go:14-124 is a runtime/pprof wrapper,
 *   - device/devices.go:74 hard-codes Health = Healthy,
 *   - metrics/metrics.go:1 is an empty package, and there is no *_test.go.
 * There is therefore no reference golden vector to check against and no Go
 * toolchain here to run one.  What this file restates is the spec frozen in
 * SURVEY.md §8(c); it is pinned by the closed-form known answers in that
 * section (tests/test_oracle.py) and by a brute-force pure-Python loop on
 * small N, nothing more.
 *
 * Spec: words are uint64 little-endian, i is the word index
 * from the buffer base, every sum is mod 2^64 (so any reduction order gives
 * the same bits).
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#define ORACLE_EXPORT __attribute__((visibility("default")))

typedef struct gp_pattern_result {
  uint64_t sum, xor_, wsum, mismatches, first_bad_index, first_bad_xor;
} gp_pattern_result;

enum { PAT_WALKING_ONES = 0, PAT_WALKING_ZEROS = 1, PAT_ADDRESS = 2 };

/* walking-ones P = 1 << ((i + p) & 63); walking-zeros its
 * complement; address-in-address pass 0 P = 8 i, pass 1 P = ~(8 i). */
static inline uint64_t pattern_word(int kind, int pass, uint64_t i) {
  switch (kind) {
    case PAT_WALKING_ONES: return 1ull << ((i + (uint64_t)pass) & 63);
    case PAT_WALKING_ZEROS: return ~(1ull << ((i + (uint64_t)pass) & 63));
    default: return (pass & 1) ? ~(8ull * i) : 8ull * i;
  }
}

ORACLE_EXPORT uint64_t gp_cpu_pattern_word(int kind, int pass, uint64_t i) {
  return pattern_word(kind, pass, i);
}

static int valid_kind_pass(int kind, int pass) {
  if (kind == PAT_WALKING_ONES || kind == PAT_WALKING_ZEROS) return pass >= 0 && pass < 64;
  if (kind == PAT_ADDRESS) return pass == 0 || pass == 1;
  return 0;
}

typedef struct {
  uint64_t* buf;
  uint64_t begin, end;
  int kind, pass, do_fill, do_verify;
  gp_pattern_result r;
} job_t;

static void* pattern_job(void* p) {
  job_t* j = (job_t*)p;
  if (j->do_fill)
    for (uint64_t i = j->begin; i < j->end; ++i) j->buf[i] = pattern_word(j->kind, j->pass, i);
  gp_pattern_result r = {0, 0, 0, 0, UINT64_MAX, 0};
  if (j->do_verify) {
    for (uint64_t i = j->begin; i < j->end; ++i) {
      uint64_t w = j->buf[i], e = pattern_word(j->kind, j->pass, i);
      r.sum += w;
      r.xor_ ^= w;
      r.wsum += w * (2 * i + 1);
      if (w != e) {
        if (r.mismatches == 0) { r.first_bad_index = i; r.first_bad_xor = w ^ e; }
        ++r.mismatches;
      }
    }
  }
  j->r = r;
  return NULL;
}

static int run_pattern(uint64_t* buf, uint64_t nwords, int kind, int pass, int nthreads,
                       int do_fill, int do_verify, gp_pattern_result* out) {
  if (!buf && nwords) return -1;
  if (!valid_kind_pass(kind, pass)) return -1;
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 1024) nthreads = 1024;
  job_t* jobs = (job_t*)calloc((size_t)nthreads, sizeof(job_t));
  pthread_t* th = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
  if (!jobs || !th) { free(jobs); free(th); return -5; }
  uint64_t chunk = (nwords + (uint64_t)nthreads - 1) / (uint64_t)nthreads;
  for (int t = 0; t < nthreads; ++t) {
    uint64_t b = chunk * (uint64_t)t, e = b + chunk;
    if (b > nwords) b = nwords;
    if (e > nwords) e = nwords;
    jobs[t] = (job_t){buf, b, e, kind, pass, do_fill, do_verify, {0, 0, 0, 0, UINT64_MAX, 0}};
    if (t > 0) pthread_create(&th[t], NULL, pattern_job, &jobs[t]);
  }
  pattern_job(&jobs[0]);
  gp_pattern_result r = jobs[0].r;
  for (int t = 1; t < nthreads; ++t) {
    pthread_join(th[t], NULL);
    const gp_pattern_result* p = &jobs[t].r;
    r.sum += p->sum;
    r.xor_ ^= p->xor_;
    r.wsum += p->wsum;
    /* chunks are in ascending index order, so the first chunk with a mismatch wins */
    if (r.mismatches == 0 && p->mismatches != 0) {
      r.first_bad_index = p->first_bad_index;
      r.first_bad_xor = p->first_bad_xor;
    }
    r.mismatches += p->mismatches;
  }
  if (out) *out = r;
  free(jobs);
  free(th);
  return 0;
}

/* fill then verify, the "Reference Go CPU probe" row of BASELINE.json configs[0]
 * (which exists only as this restatement — see the header). */
ORACLE_EXPORT int gp_cpu_pattern(uint64_t* buf, uint64_t nwords, int kind, int pass, int nthreads,
                                 gp_pattern_result* out) {
  return run_pattern(buf, nwords, kind, pass, nthreads, 1, 1, out);
}
ORACLE_EXPORT int gp_cpu_pattern_fill(uint64_t* buf, uint64_t nwords, int kind, int pass,
                                      int nthreads) {
  return run_pattern(buf, nwords, kind, pass, nthreads, 1, 0, NULL);
}
ORACLE_EXPORT int gp_cpu_pattern_verify(const uint64_t* buf, uint64_t nwords, int kind, int pass,
                                        int nthreads, gp_pattern_result* out) {
  return run_pattern((uint64_t*)buf, nwords, kind, pass, nthreads, 0, 1, out);
}

/* checksum of checksums used by gp_device_result.pattern_digest */
static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
ORACLE_EXPORT uint64_t gp_cpu_digest_step(uint64_t digest, const gp_pattern_result* r,
                                          uint32_t pass_no) {
  return digest + (r->sum ^ rotl64(r->xor_, 1) ^ rotl64(r->wsum, 2)) * (uint64_t)(pass_no + 1);
}

/* ---- counter-based generator (SURVEY.md §8c "GEMM oracle": splitmix64(seed, idx)
 *      -> U(-1,1) -> round-to-nearest-even bf16) ---------------------------- */
static inline uint64_t mix64(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + (idx + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
/* 24 random bits -> (u - 2^23) / 2^23 in [-1, 1): exact in fp32 */
static inline float gen_f32(uint64_t seed, uint64_t idx) {
  int32_t u = (int32_t)(mix64(seed, idx) >> 40);
  return (float)(u - (1 << 23)) * (1.0f / 8388608.0f);
}
static inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  if ((x & 0x7F800000u) == 0x7F800000u && (x & 0x007FFFFFu)) return (uint16_t)((x >> 16) | 0x40);
  x += 0x7FFFu + ((x >> 16) & 1u);
  return (uint16_t)(x >> 16);
}
static inline float bf16_to_f32(uint16_t h) {
  uint32_t x = (uint32_t)h << 16;
  float f;
  memcpy(&f, &x, 4);
  return f;
}

#define SEED_B_XOR 0xA5A5A5A5A5A5A5A5ull      /* GEMM operand B stream     */
#define SEED_TRIAD_B_XOR 0x1111111111111111ull /* triad operand b stream    */
#define SEED_TRIAD_C_XOR 0x2222222222222222ull /* triad operand c stream    */

ORACLE_EXPORT uint64_t gp_cpu_mix64(uint64_t seed, uint64_t idx) { return mix64(seed, idx); }
ORACLE_EXPORT float gp_cpu_gen_f32(uint64_t seed, uint64_t idx) { return gen_f32(seed, idx); }
ORACLE_EXPORT uint16_t gp_cpu_gen_bf16(uint64_t seed, uint64_t idx) {
  return f32_to_bf16_rne(gen_f32(seed, idx));
}
ORACLE_EXPORT void gp_cpu_gen_bf16_array(uint64_t seed, uint64_t idx0, uint64_t count,
                                         uint16_t* out) {
  for (uint64_t i = 0; i < count; ++i) out[i] = f32_to_bf16_rne(gen_f32(seed, idx0 + i));
}
ORACLE_EXPORT void gp_cpu_gen_f32_array(uint64_t seed, uint64_t idx0, uint64_t count, float* out) {
  for (uint64_t i = 0; i < count; ++i) out[i] = gen_f32(seed, idx0 + i);
}

/* ---- HBM stream probes (copy = memcmp-equivalent checksum;
 *      triad a = b + s*c with a single fused multiply-add, fmaf) ----------- */
ORACLE_EXPORT void gp_cpu_triad(float* a, const float* b, const float* c, float s, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) a[i] = fmaf(s, c[i], b[i]);
}
/* operands as the device generates them */
ORACLE_EXPORT void gp_cpu_triad_gen(float* a, uint64_t seed, float s, uint64_t idx0, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i)
    a[i] = fmaf(s, gen_f32(seed ^ SEED_TRIAD_C_XOR, idx0 + i), gen_f32(seed ^ SEED_TRIAD_B_XOR, idx0 + i));
}
ORACLE_EXPORT uint64_t gp_cpu_sum_words(const uint64_t* w, uint64_t n) {
  uint64_t s = 0;
  for (uint64_t i = 0; i < n; ++i) s += w[i];
  return s;
}

/* ---- GEMM: C[i][j] = SUM_k A[i][k] * B[j][k]; A is m x k, B is n x k, both
 *      K-major bf16; reference in fp64 of the bf16-rounded inputs --------- */
ORACLE_EXPORT void gp_cpu_gemm_ref(const uint16_t* a, const uint16_t* b, double* c, double* absdot,
                                   uint32_t m, uint32_t n, uint32_t k) {
  for (uint32_t i = 0; i < m; ++i)
    for (uint32_t j = 0; j < n; ++j) {
      double acc = 0.0, aacc = 0.0;
      for (uint32_t t = 0; t < k; ++t) {
        double p = (double)bf16_to_f32(a[(size_t)i * k + t]) * (double)bf16_to_f32(b[(size_t)j * k + t]);
        acc += p;
        aacc += fabs(p);
      }
      c[(size_t)i * n + j] = acc;
      if (absdot) absdot[(size_t)i * n + j] = aacc;
    }
}
/* one entry of the generated problem: returns the fp64 dot product and SUM|a||b| */
ORACLE_EXPORT double gp_cpu_gemm_sample(uint64_t seed, uint32_t k, uint32_t i, uint32_t j,
                                        double* absdot) {
  double acc = 0.0, aacc = 0.0;
  for (uint32_t t = 0; t < k; ++t) {
    double x = (double)bf16_to_f32(f32_to_bf16_rne(gen_f32(seed, (uint64_t)i * k + t)));
    double y = (double)bf16_to_f32(f32_to_bf16_rne(gen_f32(seed ^ SEED_B_XOR, (uint64_t)j * k + t)));
    acc += x * y;
    aacc += fabs(x * y);
  }
  if (absdot) *absdot = aacc;
  return acc;
}

/* ---- the bench "step" on host memory: the same seven operations gp_hbm_step
 *      enqueues on the device (include/gpuprobe.h), threads over index ranges.
 *      This is what bench.py times as cpu_baseline / --impl reference: a CPU
 *      port of the spec, NOT reference code (the reference has none). -------- */
typedef struct {
  int op; /* 0 fill, 1 copy, 2 verify, 3 triad, 4 gen */
  uint64_t begin, end;
  uint64_t* x; const uint64_t* src; float* a; const float* b; const float* c;
  int kind, pass; uint64_t seed; float s;
  gp_pattern_result r;
} step_job_t;

static void* step_job(void* p) {
  step_job_t* j = (step_job_t*)p;
  switch (j->op) {
    case 0: for (uint64_t i = j->begin; i < j->end; ++i) j->x[i] = pattern_word(j->kind, j->pass, i); break;
    case 1: memcpy(j->x + j->begin, j->src + j->begin, (j->end - j->begin) * 8); break;
    case 2: {
      gp_pattern_result r = {0, 0, 0, 0, UINT64_MAX, 0};
      for (uint64_t i = j->begin; i < j->end; ++i) {
        uint64_t w = j->src[i];
        r.sum += w; r.xor_ ^= w; r.wsum += w * (2 * i + 1);
        if (j->kind >= 0) {
          uint64_t e = pattern_word(j->kind, j->pass, i);
          if (w != e) { if (!r.mismatches) { r.first_bad_index = i; r.first_bad_xor = w ^ e; } ++r.mismatches; }
        }
      }
      j->r = r;
      break;
    }
    case 3: for (uint64_t i = j->begin; i < j->end; ++i) j->a[i] = fmaf(j->s, j->c[i], j->b[i]); break;
    case 4: for (uint64_t i = j->begin; i < j->end; ++i) j->a[i] = gen_f32(j->seed, i); break;
  }
  return NULL;
}

static void run_step_op(step_job_t proto, uint64_t n, int nthreads, gp_pattern_result* out) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 1024) nthreads = 1024;
  step_job_t* jobs = (step_job_t*)calloc((size_t)nthreads, sizeof(step_job_t));
  pthread_t* th = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
  uint64_t chunk = (n + (uint64_t)nthreads - 1) / (uint64_t)nthreads;
  for (int t = 0; t < nthreads; ++t) {
    jobs[t] = proto;
    jobs[t].begin = chunk * (uint64_t)t > n ? n : chunk * (uint64_t)t;
    jobs[t].end = jobs[t].begin + chunk > n ? n : jobs[t].begin + chunk;
    if (t > 0) pthread_create(&th[t], NULL, step_job, &jobs[t]);
  }
  step_job(&jobs[0]);
  gp_pattern_result r = jobs[0].r;
  for (int t = 1; t < nthreads; ++t) {
    pthread_join(th[t], NULL);
    if (proto.op == 2) {
      const gp_pattern_result* q = &jobs[t].r;
      r.sum += q->sum; r.xor_ ^= q->xor_; r.wsum += q->wsum;
      if (r.mismatches == 0 && q->mismatches) { r.first_bad_index = q->first_bad_index; r.first_bad_xor = q->first_bad_xor; }
      r.mismatches += q->mismatches;
    }
  }
  if (out && proto.op == 2) *out = r;
  free(jobs);
  free(th);
}

/* arena: 5 * bytes of host memory (X, Y, Z, b, c).  gen != 0 regenerates b and c. */
ORACLE_EXPORT int gp_cpu_hbm_step(void* arena, uint64_t bytes, uint32_t step, uint64_t seed,
                                  int gen, int nthreads, gp_pattern_result* out /*[3]*/) {
  if (!arena || bytes == 0 || (bytes % 512)) return -1;
  uint64_t nw = bytes / 8, nf = bytes / 4;
  uint64_t* X = (uint64_t*)arena;
  uint64_t* Y = (uint64_t*)((char*)arena + bytes);
  float* Z = (float*)((char*)arena + 2 * bytes);
  float* B = (float*)((char*)arena + 3 * bytes);
  float* Cc = (float*)((char*)arena + 4 * bytes);
  step_job_t j;
  memset(&j, 0, sizeof(j));
  if (gen) {
    j.op = 4; j.a = B; j.seed = seed ^ SEED_TRIAD_B_XOR; run_step_op(j, nf, nthreads, NULL);
    j.op = 4; j.a = Cc; j.seed = seed ^ SEED_TRIAD_C_XOR; run_step_op(j, nf, nthreads, NULL);
  }
  int wo = (int)(step & 63), ad = (int)(step & 1);
  j.op = 0; j.x = X; j.kind = PAT_WALKING_ONES; j.pass = wo; run_step_op(j, nw, nthreads, NULL);
  j.op = 1; j.x = Y; j.src = X; run_step_op(j, nw, nthreads, NULL);
  j.op = 2; j.src = Y; j.kind = PAT_WALKING_ONES; j.pass = wo; run_step_op(j, nw, nthreads, &out[0]);
  j.op = 0; j.x = X; j.kind = PAT_ADDRESS; j.pass = ad; run_step_op(j, nw, nthreads, NULL);
  j.op = 2; j.src = X; j.kind = PAT_ADDRESS; j.pass = ad; run_step_op(j, nw, nthreads, &out[1]);
  j.op = 3; j.a = Z; j.b = B; j.c = Cc; j.s = 3.0f; run_step_op(j, nf, nthreads, NULL);
  j.op = 2; j.src = (const uint64_t*)Z; j.kind = -1; j.pass = 0; run_step_op(j, nw, nthreads, &out[2]);
  return 0;
}

/* Known answer for the triad probe: SUM and XOR (mod 2^64) of the u64 words of a[i] = fmaf(s, c[i], b[i]) with b, c
 * from the counter generator — what the device leaves in `a`, folded exactly as the verify kernel folds it.  Streaming
 * (nothing is stored), so the 16 GiB default configuration can be computed on any host.  The device grades its own
 * triad with the same fmaf on the same chip; comparing against THIS number is what would catch a broken FMA. */
typedef struct { uint64_t seed, i0, i1; float s; uint64_t sum, xr; } triad_ck_job;
static void* triad_ck_run(void* p) {
  triad_ck_job* j = (triad_ck_job*)p;
  uint64_t sum = 0, xr = 0;
  for (uint64_t w = j->i0; w < j->i1; ++w) {
    uint32_t lo, hi;
    const float a0 = fmaf(j->s, gen_f32(j->seed ^ 0x2222222222222222ull, 2 * w), gen_f32(j->seed ^ 0x1111111111111111ull, 2 * w));
    const float a1 = fmaf(j->s, gen_f32(j->seed ^ 0x2222222222222222ull, 2 * w + 1), gen_f32(j->seed ^ 0x1111111111111111ull, 2 * w + 1));
    memcpy(&lo, &a0, 4);
    memcpy(&hi, &a1, 4);
    const uint64_t word = (uint64_t)lo | ((uint64_t)hi << 32);
    sum += word;
    xr ^= word;
  }
  j->sum = sum;
  j->xr = xr;
  return NULL;
}
ORACLE_EXPORT void gp_cpu_triad_checksum(uint64_t seed, uint64_t nfloats, float s, int nthreads, uint64_t* sum, uint64_t* xr) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  const uint64_t nwords = nfloats / 2;
  pthread_t th[256];
  triad_ck_job jobs[256];
  for (int t = 0; t < nthreads; ++t) {
    jobs[t].seed = seed; jobs[t].s = s;
    jobs[t].i0 = nwords * (uint64_t)t / (uint64_t)nthreads;
    jobs[t].i1 = nwords * (uint64_t)(t + 1) / (uint64_t)nthreads;
    pthread_create(&th[t], NULL, triad_ck_run, &jobs[t]);
  }
  uint64_t S = 0, X = 0;
  for (int t = 0; t < nthreads; ++t) { pthread_join(th[t], NULL); S += jobs[t].sum; X ^= jobs[t].xr; }
  *sum = S;
  *xr = X;
}
