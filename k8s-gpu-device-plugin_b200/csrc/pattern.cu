// pattern.cu — integer memory-pattern fill and verify (K2a / K2b of SURVEY.md §8).
//
// No reference counterpart exists (device/devices.go:74 sets Health without
// looking at the device); the arithmetic is the spec of SURVEY.md §8(c) and
// must be bit-exact against oracle/gpuprobe_oracle.c.
//
// Both kernels are HBM-bound integer work: 256-bit accesses (LDG/STG.E.256),
// one CTA per contiguous 32 KiB chunk.
//   fill   : N bytes written, nothing read.
//   verify : N bytes read; sum / xor / position-weighted sum (all mod 2^64, so
//            reduction order is irrelevant), mismatch count, first bad index and
//            its xor.  Uses kVerifyScratchBytes of scratch (accumulator slots),
//            left in its rest state by every launch.
#include "common.cuh"
#include "kernels.h"

namespace gp {

// Launch shape, measured with tools/hbm_sweep.cu on B200 (profiles/hbm_sweep_r01.txt):
// a NON-persistent grid of one CTA per contiguous 32 KiB chunk lets the hardware CTA
// dispatcher sweep a tight, linearly advancing window over DRAM: 7.60 TB/s fill and
// 7.59 TB/s read, against 6.35 / 7.38 TB/s for a persistent grid-stride grid of the same
// instructions.  Every thread issues its four 256-bit accesses up front, no loop.
constexpr int kThreads = 256;
constexpr int kVecPerThread = 4;
constexpr uint64_t kChunkVec = (uint64_t)kThreads * kVecPerThread;  // 1024 x 32 B = 32 KiB per CTA

template <int KIND>
__global__ void __launch_bounds__(kThreads)
pattern_fill_kernel(uint64_t* __restrict__ words, uint64_t nvec, int pass) {
  const uint64_t g0 = (uint64_t)blockIdx.x * kChunkVec + threadIdx.x;
#pragma unroll
  for (int u = 0; u < kVecPerThread; ++u) {
    const uint64_t g = g0 + (uint64_t)u * kThreads;
    if (g < nvec) {
      const uint64_t i = g << 2;
      u64x4 v;
#pragma unroll
      for (int j = 0; j < 4; ++j) v.v[j] = pattern_word<KIND>(pass, i + j);
      st256_stream(words + i, v);
    }
  }
}

struct Acc {
  uint64_t sum, xr, wsum, mism, first_idx, first_xor;  // first_xor = w ^ P at first_idx, captured when it was seen
};

__device__ __forceinline__ void acc_init(Acc& a) {
  a.sum = 0; a.xr = 0; a.wsum = 0; a.mism = 0; a.first_idx = ~0ull; a.first_xor = 0;
}

// {index, xor} pair <- mine if my index is lower: a 128-bit compare-and-swap loop (cold path only: a warp
// comes here once per launch, and only if it saw a mismatch).  The xor therefore always belongs to the index
// it is stored with, and it is the value seen when the word was read — a transient fault that a later re-read
// would no longer show is still reported with the bits that were wrong.
__device__ __forceinline__ void first_bad_min(VerifySlot* s, uint64_t idx, uint64_t xr) {
  uint64_t oi = *reinterpret_cast<volatile uint64_t*>(&s->first_bad_index);
  uint64_t ox = *reinterpret_cast<volatile uint64_t*>(&s->first_bad_xor);
  while (idx < oi) {
    uint64_t gi, gx;
    asm volatile(
        "{\n\t.reg .b128 c, n, o;\n\t"
        "mov.b128 c, {%3, %4};\n\t"
        "mov.b128 n, {%5, %6};\n\t"
        "atom.global.cas.b128 o, [%2], c, n;\n\t"
        "mov.b128 {%0, %1}, o;\n\t}"
        : "=l"(gi), "=l"(gx)
        : "l"(&s->first_bad_index), "l"(oi), "l"(ox), "l"(idx), "l"(xr)
        : "memory");
    if (gi == oi && gx == ox) break;
    oi = gi; ox = gx;
  }
}

// word-by-word compare of one vector; only reached when the vector is known to differ (or on
// the ragged last chunk).  Everything travels in registers so the caller's vectors never get
// a stack home: returns the lowest bad word index (all-ones if none) and adds to *count.
template <int KIND>
__device__ __noinline__ uint64_t compare_words_slow(uint64_t w0, uint64_t w1, uint64_t w2,
                                                    uint64_t w3, uint64_t i, int pass,
                                                    uint32_t& count, uint64_t& first_xor) {
  uint64_t first = ~0ull;
  const uint64_t w[4] = {w0, w1, w2, w3};
#pragma unroll
  for (int j = 3; j >= 0; --j) {
    const uint64_t x = w[j] ^ pattern_word<KIND>(pass, i + j);
    if (x) { first = i + j; first_xor = x; ++count; }
  }
  return first;
}
template <int KIND>
__device__ __forceinline__ void compare_vec_slow(Acc& a, const u64x4& v, uint64_t g, int pass) {
  uint32_t count = 0;
  uint64_t fx = 0;
  const uint64_t first = compare_words_slow<KIND>(v.v[0], v.v[1], v.v[2], v.v[3], g << 2, pass, count, fx);
  if (first < a.first_idx) { a.first_idx = first; a.first_xor = fx; }
  a.mism += count;
}

// generic single-vector accumulate (ragged last chunk only)
template <int KIND>
__device__ __forceinline__ void acc_vec(Acc& a, const u64x4& v, uint64_t g, int pass) {
  const uint64_t i = g << 2;
  const uint64_t s4 = (v.v[0] + v.v[1]) + (v.v[2] + v.v[3]);
  a.sum += s4;
  a.xr ^= (v.v[0] ^ v.v[1]) ^ (v.v[2] ^ v.v[3]);
  a.wsum += (2 * i + 1) * s4 + 2 * (v.v[1] + 2 * v.v[2] + 3 * v.v[3]);
  if constexpr (KIND != kPatNone) compare_vec_slow<KIND>(a, v, g, pass);
}

// Reduce across the CTA (finalize kernel only); the result is valid in thread 0.
// lowest first_idx over the warp, with the xor that belongs to it
__device__ __forceinline__ void warp_min_first(Acc& a) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) {
    const uint64_t oi = shfl_xor_u64(a.first_idx, m);
    const uint64_t ox = shfl_xor_u64(a.first_xor, m);
    if (oi < a.first_idx) { a.first_idx = oi; a.first_xor = ox; }
  }
}
__device__ __forceinline__ void block_reduce(Acc& a) {
  __shared__ uint64_t sh[kThreads / 32][6];
  a.sum = warp_sum_u64(a.sum);
  a.xr = warp_xor_u64(a.xr);
  a.wsum = warp_sum_u64(a.wsum);
  a.mism = warp_sum_u64(a.mism);
  warp_min_first(a);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    sh[warp][0] = a.sum; sh[warp][1] = a.xr; sh[warp][2] = a.wsum;
    sh[warp][3] = a.mism; sh[warp][4] = a.first_idx; sh[warp][5] = a.first_xor;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < kThreads / 32; ++w) {
      a.sum += sh[w][0]; a.xr ^= sh[w][1]; a.wsum += sh[w][2]; a.mism += sh[w][3];
      if (sh[w][4] < a.first_idx) { a.first_idx = sh[w][4]; a.first_xor = sh[w][5]; }
    }
  }
}

// verify: N bytes read.  The kernel is close to issue-bound on the integer pipes (an earlier
// word-at-a-time version reached 6.1–6.5 TB/s while a bare read reaches 7.6), so the inner
// loop is built from per-thread invariants:
//   * a thread's vectors are 256 vectors = 1024 words apart, a multiple of 64 words, so its
//     four expected walking-ones words are the same for every vector: one xor/or tree and one
//     branch per 256-bit vector;
//   * address words of a vector share their upper bits: e_j = 32 g | 8 j;
//   * SUM_i w_i (2i+1) over the thread's vectors needs ONE 64-bit multiply:
//       (8 g0 + 1) S + 2048 SUM_u u s4_u + 2 (A1 + 2 A2 + 3 A3),  A_j = SUM_u w_{u,j}.
// Each warp folds its partial with shuffles and lane 0 adds it into one of kVerifySlots
// accumulator slots in scratch memory with fire-and-forget reductions (RED.ADD / RED.XOR /
// RED.MIN at L2; all arithmetic is mod 2^64, so the order is irrelevant and the result
// bit-exact).  No CTA waits on anything: a first version elected the last CTA with a returning
// ticket atomic and lost 20 % of the bandwidth to that one hot address.  A one-CTA finalize
// launch on the same stream folds the slots, looks up the xor at the first bad index, writes
// the 48-byte result with plain stores and returns the slots to their rest state — so verify
// launches chain on a stream with no memset in between.
template <int KIND>
__global__ void __launch_bounds__(kThreads)
pattern_verify_kernel(const uint64_t* __restrict__ words, uint64_t nvec, int pass,
                      VerifySlot* __restrict__ slots) {
  Acc a;
  acc_init(a);
  const uint64_t g0 = (uint64_t)blockIdx.x * kChunkVec + threadIdx.x;
  if (g0 - threadIdx.x + kChunkVec <= nvec) {  // whole chunk: every load issued up front
    u64x4 v[kVecPerThread];
#pragma unroll
    for (int u = 0; u < kVecPerThread; ++u) v[u] = ld256_stream(words + ((g0 + (uint64_t)u * kThreads) << 2));
    uint64_t R = 0, B = 0, A1 = 0, A2 = 0, A3 = 0, X = 0, S = 0;
#pragma unroll
    for (int u = kVecPerThread - 1; u >= 0; --u) {
      const uint64_t s4 = (v[u].v[0] + v[u].v[1]) + (v[u].v[2] + v[u].v[3]);
      X ^= (v[u].v[0] ^ v[u].v[1]) ^ (v[u].v[2] ^ v[u].v[3]);
      A1 += v[u].v[1]; A2 += v[u].v[2]; A3 += v[u].v[3];
      if (u > 0) { R += s4; B += R; } else { S = R + s4; }
    }
    a.sum = S;
    a.xr = X;
    a.wsum = (8 * g0 + 1) * S + (uint64_t)(8 * kThreads) * B + 2 * (A1 + 2 * A2 + 3 * A3);
    if constexpr (KIND == GP_PAT_WALKING_ONES || KIND == GP_PAT_WALKING_ZEROS) {
      uint64_t e[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = pattern_word<KIND>(pass, (g0 << 2) + j);  // same for every u
#pragma unroll
      for (int u = 0; u < kVecPerThread; ++u) {
        const uint64_t m = ((v[u].v[0] ^ e[0]) | (v[u].v[1] ^ e[1])) | ((v[u].v[2] ^ e[2]) | (v[u].v[3] ^ e[3]));
        if (m) compare_vec_slow<KIND>(a, v[u], g0 + (uint64_t)u * kThreads, pass);
      }
    } else if constexpr (KIND == GP_PAT_ADDRESS) {
      const uint64_t inv = (pass & 1) ? ~0ull : 0ull;
#pragma unroll
      for (int u = 0; u < kVecPerThread; ++u) {
        const uint64_t e0 = ((g0 + (uint64_t)u * kThreads) << 5) ^ inv;  // 8 * (4 g) ; low five bits free
        const uint64_t m = ((v[u].v[0] ^ e0) | (v[u].v[1] ^ e0 ^ 8)) | ((v[u].v[2] ^ e0 ^ 16) | (v[u].v[3] ^ e0 ^ 24));
        if (m) compare_vec_slow<KIND>(a, v[u], g0 + (uint64_t)u * kThreads, pass);
      }
    }
  } else {  // ragged last chunk
    for (int u = 0; u < kVecPerThread; ++u) {
      const uint64_t g = g0 + (uint64_t)u * kThreads;
      if (g < nvec) acc_vec<KIND>(a, ld256_stream(words + (g << 2)), g, pass);
    }
  }
  a.sum = warp_sum_u64(a.sum);
  a.xr = warp_xor_u64(a.xr);
  a.wsum = warp_sum_u64(a.wsum);
  const bool any_bad = __any_sync(0xffffffffu, a.mism != 0);
  if (any_bad) {  // cold
    a.mism = warp_sum_u64(a.mism);
    warp_min_first(a);
  }
  if ((threadIdx.x & 31) == 0) {
    VerifySlot* s = slots + ((blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5)) & (kVerifySlots - 1));
    atomicAdd(reinterpret_cast<unsigned long long*>(&s->sum), (unsigned long long)a.sum);
    atomicXor(reinterpret_cast<unsigned long long*>(&s->xor_), (unsigned long long)a.xr);
    atomicAdd(reinterpret_cast<unsigned long long*>(&s->wsum), (unsigned long long)a.wsum);
    if (any_bad) {
      atomicAdd(reinterpret_cast<unsigned long long*>(&s->mismatches), (unsigned long long)a.mism);
      first_bad_min(s, a.first_idx, a.first_xor);
    }
  }
}

template <int KIND>
__global__ void __launch_bounds__(kThreads)
pattern_verify_finalize_kernel(const uint64_t* __restrict__ words, int pass,
                               VerifySlot* __restrict__ slots,
                               gp_pattern_result* __restrict__ result) {
  Acc a;
  acc_init(a);
  for (int b = threadIdx.x; b < kVerifySlots; b += kThreads) {
    VerifySlot* s = slots + b;
    a.sum += s->sum; a.xr ^= s->xor_; a.wsum += s->wsum; a.mism += s->mismatches;
    if (s->first_bad_index < a.first_idx) { a.first_idx = s->first_bad_index; a.first_xor = s->first_bad_xor; }
    s->sum = 0; s->xor_ = 0; s->wsum = 0; s->mismatches = 0; s->first_bad_index = ~0ull; s->first_bad_xor = 0;  // rest state
  }
  block_reduce(a);
  if (threadIdx.x == 0) {
    gp_pattern_result r;
    r.sum = a.sum; r.xor_ = a.xr; r.wsum = a.wsum; r.mismatches = a.mism;
    r.first_bad_index = a.first_idx;
    r.first_bad_xor = a.first_idx != ~0ull ? a.first_xor : 0;  // the bits seen when the word was read, not a re-read
    *result = r;
  }
}

__global__ void verify_scratch_init_kernel(VerifySlot* slots) {
  for (int b = threadIdx.x; b < kVerifySlots; b += blockDim.x) {
    slots[b].sum = 0; slots[b].xor_ = 0; slots[b].wsum = 0; slots[b].mismatches = 0;
    slots[b].first_bad_index = ~0ull; slots[b].first_bad_xor = 0; slots[b].pad[0] = slots[b].pad[1] = 0;
  }
}

cudaError_t launch_verify_scratch_init(void* scratch, cudaStream_t stream) {
  verify_scratch_init_kernel<<<1, 256, 0, stream>>>(reinterpret_cast<VerifySlot*>(scratch));
  return cudaGetLastError();
}

// The ABI requires nwords % 64 == 0, so nwords is always a whole number of
// 256-bit vectors; the engine checks it.
static unsigned int grid_for(uint64_t nvec) {
  const uint64_t g = (nvec + kChunkVec - 1) / kChunkVec;
  return (unsigned int)(g < 1 ? 1 : g);
}

cudaError_t launch_pattern_fill(uint64_t* words, uint64_t nwords, int kind, int pass, int sm_count,
                                cudaStream_t stream) {
  (void)sm_count;
  const uint64_t nvec = nwords >> 2;
  if (nvec == 0) return cudaSuccess;
  const unsigned int grid = grid_for(nvec);
  switch (kind) {
    case GP_PAT_WALKING_ONES:
      pattern_fill_kernel<GP_PAT_WALKING_ONES><<<grid, kThreads, 0, stream>>>(words, nvec, pass);
      break;
    case GP_PAT_WALKING_ZEROS:
      pattern_fill_kernel<GP_PAT_WALKING_ZEROS><<<grid, kThreads, 0, stream>>>(words, nvec, pass);
      break;
    case GP_PAT_ADDRESS:
      pattern_fill_kernel<GP_PAT_ADDRESS><<<grid, kThreads, 0, stream>>>(words, nvec, pass);
      break;
    default:
      return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_pattern_verify(const uint64_t* words, uint64_t nwords, int kind, int pass,
                                  void* scratch, gp_pattern_result* d_result, int sm_count,
                                  cudaStream_t stream) {
  (void)sm_count;
  const uint64_t nvec = nwords >> 2;
  const unsigned int grid = grid_for(nvec);
  VerifySlot* slots = reinterpret_cast<VerifySlot*>(scratch);
#define GP_VERIFY_CASE(K)                                                                        \
  case K:                                                                                        \
    if (nvec) pattern_verify_kernel<K><<<grid, kThreads, 0, stream>>>(words, nvec, pass, slots); \
    pattern_verify_finalize_kernel<K><<<1, kThreads, 0, stream>>>(words, pass, slots, d_result); \
    break;
  switch (kind) {
    GP_VERIFY_CASE(GP_PAT_WALKING_ONES)
    GP_VERIFY_CASE(GP_PAT_WALKING_ZEROS)
    GP_VERIFY_CASE(GP_PAT_ADDRESS)
    GP_VERIFY_CASE(kPatNone)
#undef GP_VERIFY_CASE
    default:
      return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_verify_fold(void* scratch, gp_pattern_result* d_result, cudaStream_t stream) {
  pattern_verify_finalize_kernel<kPatNone><<<1, kThreads, 0, stream>>>(nullptr, 0, reinterpret_cast<VerifySlot*>(scratch), d_result);
  return cudaGetLastError();
}

__global__ void bitflip_kernel(uint64_t* words, uint64_t word_index, int bit) {
  words[word_index] ^= (1ull << (bit & 63));
}

cudaError_t launch_bitflip(uint64_t* words, uint64_t word_index, int bit, cudaStream_t stream) {
  bitflip_kernel<<<1, 1, 0, stream>>>(words, word_index, bit);
  return cudaGetLastError();
}

}  // namespace gp
