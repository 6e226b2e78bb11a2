// common.cuh — shared device helpers for the probe kernels (sm_100a only).
//
// The arithmetic here is the spec frozen in SURVEY.md §8(c); the CPU twin is
// oracle/gpuprobe_oracle.c (test infrastructure, never linked here).  The
// reference has no counterpart: device/devices.go:74 hard-codes Healthy.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/gpuprobe.h"

namespace gp {

// ---- pattern words ----------------------------------------
// kind/pass are warp-uniform; i is the word index from the arena base.
template <int KIND>
__device__ __forceinline__ uint64_t pattern_word(int pass, uint64_t i) {
  if constexpr (KIND == GP_PAT_WALKING_ONES) {
    return 1ull << ((unsigned)(i + (uint64_t)pass) & 63u);
  } else if constexpr (KIND == GP_PAT_WALKING_ZEROS) {
    return ~(1ull << ((unsigned)(i + (uint64_t)pass) & 63u));
  } else {
    const uint64_t a = i << 3;
    return (pass & 1) ? ~a : a;
  }
}

// ---- counter-based generator: splitmix64 finaliser over (seed, idx) -------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + (idx + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// 24 random bits -> (u - 2^23) / 2^23 in [-1, 1): exactly representable in fp32
__host__ __device__ __forceinline__ float gen_f32(uint64_t seed, uint64_t idx) {
  const int32_t u = (int32_t)(mix64(seed, idx) >> 40);
  return (float)(u - (1 << 23)) * (1.0f / 8388608.0f);
}
__host__ __device__ __forceinline__ uint16_t f32_to_bf16_rne(float f) {
#ifdef __CUDA_ARCH__
  uint32_t x = __float_as_uint(f);
#else
  uint32_t x;
  memcpy(&x, &f, 4);
#endif
  x += 0x7FFFu + ((x >> 16) & 1u);  // generator output is finite: no NaN branch needed
  return (uint16_t)(x >> 16);
}
__host__ __device__ __forceinline__ float bf16_bits_to_f32(uint16_t h) {
#ifdef __CUDA_ARCH__
  return __uint_as_float((uint32_t)h << 16);
#else
  uint32_t x = (uint32_t)h << 16;
  float f;
  memcpy(&f, &x, 4);
  return f;
#endif
}

constexpr uint64_t kSeedBXor = 0xA5A5A5A5A5A5A5A5ull;       // GEMM operand B stream
constexpr uint64_t kSeedTriadBXor = 0x1111111111111111ull;  // triad operand b stream
constexpr uint64_t kSeedTriadCXor = 0x2222222222222222ull;  // triad operand c stream

// ---- 256-bit global accesses (LDG.E.256 / STG.E.256 on sm_100) ------------
struct alignas(32) u64x4 {
  uint64_t v[4];
};

__device__ __forceinline__ u64x4 ld256_stream(const void* p) {
  u64x4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
               : "=l"(r.v[0]), "=l"(r.v[1]), "=l"(r.v[2]), "=l"(r.v[3])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st256_stream(void* p, const u64x4& r) {
  asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(r.v[0]),
               "l"(r.v[1]), "l"(r.v[2]), "l"(r.v[3])
               : "memory");
}

// ---- warp reductions over u64 ---------------------------------------------
__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int m) {
  return __shfl_xor_sync(0xffffffffu, v, m);
}
__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v += shfl_xor_u64(v, m);
  return v;
}
__device__ __forceinline__ uint64_t warp_xor_u64(uint64_t v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v ^= shfl_xor_u64(v, m);
  return v;
}

// ---- mbarrier / bulk-copy (TMA 1-D) PTX ------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t bar_smem_addr, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(bar_smem_addr),
      "r"(parity)
      : "memory");
}
// global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// shared -> global, tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// order generic-proxy smem writes before async-proxy (TMA) reads of the same bytes
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

}  // namespace gp
