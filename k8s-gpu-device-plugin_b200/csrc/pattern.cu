// pattern.cu — integer memory-pattern fill and verify (K2a / K2b of SURVEY.md §8).
//
// No reference counterpart exists (device/devices.go:74 sets Health without
// looking at the device); the arithmetic is the spec of SURVEY.md §8(c) and
// must be bit-exact against oracle/gpuprobe_oracle.c.
//
// Both kernels are HBM-bound integer work: 256-bit accesses, a persistent grid
// of sm_count * kBlocksPerSm CTAs walking the arena with a grid stride so that
// neighbouring CTAs touch neighbouring DRAM pages at the same time.
//   fill   : N bytes written, nothing read.
//   verify : N bytes read; sum / xor / position-weighted sum (all mod 2^64, so
//            reduction order is irrelevant), mismatch count, first bad index and
//            its xor.  Warp-shuffle reduction -> one partial per CTA in scratch
//            -> the last CTA to finish (ticket counter) folds the partials and
//            writes the 48-byte result with plain stores.  The ticket returns to
//            zero, so back-to-back launches need no memset and no atomics on
//            the result.
#include "common.cuh"
#include "kernels.h"

namespace gp {

constexpr int kThreads = 256;

// Persistent grids: exactly as many CTAs as fit at once (one wave), asked of the
// occupancy calculator per kernel so a register-count change cannot create a tail wave.
template <auto kernel>
static int resident_blocks_per_sm() {
  static int cached = 0;
  if (cached == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) n = 1;
    cached = n;
  }
  return cached;
}

template <int KIND>
__global__ void __launch_bounds__(kThreads)
pattern_fill_kernel(uint64_t* __restrict__ words, uint64_t nvec, int pass) {
  const uint64_t stride = (uint64_t)gridDim.x * kThreads;
  for (uint64_t g = (uint64_t)blockIdx.x * kThreads + threadIdx.x; g < nvec; g += stride) {
    const uint64_t i = g << 2;
    u64x4 v;
#pragma unroll
    for (int j = 0; j < 4; ++j) v.v[j] = pattern_word<KIND>(pass, i + j);
    st256_stream(words + i, v);
  }
}

struct Acc {
  uint64_t sum, xr, wsum, mism, first_idx, first_xor;
};

__device__ __forceinline__ void acc_init(Acc& a) {
  a.sum = 0; a.xr = 0; a.wsum = 0; a.mism = 0; a.first_idx = ~0ull; a.first_xor = 0;
}

template <int KIND>
__device__ __forceinline__ void acc_vec(Acc& a, const u64x4& v, uint64_t g, int pass) {
  const uint64_t i = g << 2;
  const uint64_t s4 = (v.v[0] + v.v[1]) + (v.v[2] + v.v[3]);
  a.sum += s4;
  a.xr ^= (v.v[0] ^ v.v[1]) ^ (v.v[2] ^ v.v[3]);
  // SUM_j w_j (2(i+j)+1) = (2i+1) s4 + 2 (w1 + 2 w2 + 3 w3): one 64-bit multiply per vector
  a.wsum += (2 * i + 1) * s4 + 2 * (v.v[1] + 2 * v.v[2] + 3 * v.v[3]);
  if constexpr (KIND != kPatNone) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint64_t e = pattern_word<KIND>(pass, i + j);
      if (v.v[j] != e) {  // cold path
        if (a.mism == 0) { a.first_idx = i + j; a.first_xor = v.v[j] ^ e; }
        ++a.mism;
      }
    }
  }
}

__device__ __forceinline__ void acc_merge_first(Acc& a, uint64_t idx, uint64_t x) {
  if (idx < a.first_idx) { a.first_idx = idx; a.first_xor = x; }
}

// Reduce across the CTA; the result is valid in thread 0.
__device__ __forceinline__ void block_reduce(Acc& a) {
  __shared__ uint64_t sh[kThreads / 32][6];
  a.sum = warp_sum_u64(a.sum);
  a.xr = warp_xor_u64(a.xr);
  a.wsum = warp_sum_u64(a.wsum);
  a.mism = warp_sum_u64(a.mism);
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) {
    const uint64_t oi = shfl_xor_u64(a.first_idx, m);
    const uint64_t ox = shfl_xor_u64(a.first_xor, m);
    acc_merge_first(a, oi, ox);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();  // sh may still be read by a previous call
  if (lane == 0) {
    sh[warp][0] = a.sum; sh[warp][1] = a.xr; sh[warp][2] = a.wsum;
    sh[warp][3] = a.mism; sh[warp][4] = a.first_idx; sh[warp][5] = a.first_xor;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < kThreads / 32; ++w) {
      a.sum += sh[w][0]; a.xr ^= sh[w][1]; a.wsum += sh[w][2]; a.mism += sh[w][3];
      acc_merge_first(a, sh[w][4], sh[w][5]);
    }
  }
}

template <int KIND>
__global__ void __launch_bounds__(kThreads, 4)
pattern_verify_kernel(const uint64_t* __restrict__ words, uint64_t nvec, int pass,
                      VerifyPartial* __restrict__ partials, unsigned int* __restrict__ ticket,
                      gp_pattern_result* __restrict__ result) {
  Acc a;
  acc_init(a);
  const uint64_t stride = (uint64_t)gridDim.x * kThreads;
  uint64_t g = (uint64_t)blockIdx.x * kThreads + threadIdx.x;
  // four independent 256-bit loads in flight per thread
  for (; g + 3 * stride < nvec; g += 4 * stride) {
    const u64x4 v0 = ld256_stream(words + (g << 2));
    const u64x4 v1 = ld256_stream(words + ((g + stride) << 2));
    const u64x4 v2 = ld256_stream(words + ((g + 2 * stride) << 2));
    const u64x4 v3 = ld256_stream(words + ((g + 3 * stride) << 2));
    acc_vec<KIND>(a, v0, g, pass);
    acc_vec<KIND>(a, v1, g + stride, pass);
    acc_vec<KIND>(a, v2, g + 2 * stride, pass);
    acc_vec<KIND>(a, v3, g + 3 * stride, pass);
  }
  for (; g < nvec; g += stride) {
    const u64x4 v = ld256_stream(words + (g << 2));
    acc_vec<KIND>(a, v, g, pass);
  }
  block_reduce(a);

  __shared__ bool is_last;
  if (threadIdx.x == 0) {
    VerifyPartial p{a.sum, a.xr, a.wsum, a.mism, a.first_idx, a.first_xor};
    partials[blockIdx.x] = p;
    __threadfence();
    const unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  acc_init(a);
  for (unsigned int b = threadIdx.x; b < gridDim.x; b += kThreads) {
    const volatile VerifyPartial* p = partials + b;
    a.sum += p->sum; a.xr ^= p->xor_; a.wsum += p->wsum; a.mism += p->mismatches;
    acc_merge_first(a, p->first_bad_index, p->first_bad_xor);
  }
  block_reduce(a);
  if (threadIdx.x == 0) {
    gp_pattern_result r;
    r.sum = a.sum; r.xor_ = a.xr; r.wsum = a.wsum; r.mismatches = a.mism;
    r.first_bad_index = a.first_idx; r.first_bad_xor = a.first_xor;
    *result = r;
    *ticket = 0;  // at rest for the next launch on this stream
  }
}

// A handful of trailing words (nwords % 4) never occurs: the ABI requires
// nwords % 64 == 0, checked in the engine.

static int grid_for(uint64_t nvec, int sm_count, int blocks_per_sm) {
  uint64_t want = (nvec + kThreads - 1) / kThreads;
  uint64_t cap = (uint64_t)sm_count * blocks_per_sm;
  if (cap > (uint64_t)kMaxVerifyBlocks) cap = kMaxVerifyBlocks;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

cudaError_t launch_pattern_fill(uint64_t* words, uint64_t nwords, int kind, int pass, int sm_count,
                                cudaStream_t stream) {
  const uint64_t nvec = nwords >> 2;
  if (nvec == 0) return cudaSuccess;
#define GP_FILL_CASE(K)                                                              \
  case K: {                                                                          \
    const int grid = grid_for(nvec, sm_count, resident_blocks_per_sm<pattern_fill_kernel<K>>()); \
    pattern_fill_kernel<K><<<grid, kThreads, 0, stream>>>(words, nvec, pass);        \
    break;                                                                           \
  }
  switch (kind) {
    GP_FILL_CASE(GP_PAT_WALKING_ONES)
    GP_FILL_CASE(GP_PAT_WALKING_ZEROS)
    GP_FILL_CASE(GP_PAT_ADDRESS)
#undef GP_FILL_CASE
    default:
      return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_pattern_verify(const uint64_t* words, uint64_t nwords, int kind, int pass,
                                  void* scratch, gp_pattern_result* d_result, int sm_count,
                                  cudaStream_t stream) {
  const uint64_t nvec = nwords >> 2;
  VerifyPartial* partials = reinterpret_cast<VerifyPartial*>(scratch);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(scratch) +
                                                         kMaxVerifyBlocks * sizeof(VerifyPartial));
#define GP_VERIFY_CASE(K)                                                              \
  case K: {                                                                            \
    const int grid = grid_for(nvec, sm_count, resident_blocks_per_sm<pattern_verify_kernel<K>>()); \
    pattern_verify_kernel<K><<<grid, kThreads, 0, stream>>>(words, nvec, pass, partials, ticket, d_result); \
    break;                                                                             \
  }
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

__global__ void bitflip_kernel(uint64_t* words, uint64_t word_index, int bit) {
  words[word_index] ^= (1ull << (bit & 63));
}

cudaError_t launch_bitflip(uint64_t* words, uint64_t word_index, int bit, cudaStream_t stream) {
  bitflip_kernel<<<1, 1, 0, stream>>>(words, word_index, bit);
  return cudaGetLastError();
}

}  // namespace gp
