// stream.cu — HBM stream probes: copy (K1a) and triad (K1b) of SURVEY.md §8.
//
// No reference counterpart (benchmark/benchmark.go:54-89 profiles the Go
// runtime, it moves no device memory).  Algorithmic bytes per launch
// (SURVEY.md §8d): copy 2N, triad 3N, N = bytes per array.
//
// Two hand-written variants of each, selected by gp_config.stream_variant, both
// launched as NON-persistent grids of one CTA per contiguous chunk (see the note
// at kCopyVecPerThread):
//   VEC : every thread issues its 256-bit loads, then its 256-bit stores
//         (LDG.E.256 / STG.E.256, L1 no-allocate).
//   TMA : cp.async.bulk global->shared (UBLKCP) completing on mbarriers and
//         cp.async.bulk shared->global; copy never touches the data with a
//         thread, triad combines the b and c tiles in place in shared memory.
// On B200 the two are within noise of each other (copy 6.93 vs 6.95 TB/s, triad
// 7.08 vs 7.08 TB/s, profiles/hbm_sweep_r01.txt); AUTO picks TMA.
#include "common.cuh"
#include "kernels.h"

namespace gp {

constexpr int kVecThreads = 256;
// Launch shape measured with tools/hbm_sweep.cu (profiles/hbm_sweep_r01.txt): one CTA per
// contiguous chunk, no loop, all loads of a thread issued before its stores.  The hardware
// CTA dispatcher then sweeps a tight window over DRAM: copy 6.95 TB/s against 6.2–6.6 TB/s
// for persistent grid-stride grids and 6.57 TB/s for cudaMemcpyAsync.
constexpr int kCopyVecPerThread = 4;  // 32 KiB per CTA
constexpr int kTriVecPerThread = 2;   // 16 KiB of each array per CTA

// ---------------------------------------------------------------- VEC copy --
__global__ void __launch_bounds__(kVecThreads)
copy_vec_kernel(const u64x4* __restrict__ src, u64x4* __restrict__ dst, uint64_t nvec) {
  const uint64_t g0 = (uint64_t)blockIdx.x * (kVecThreads * kCopyVecPerThread) + threadIdx.x;
  if (g0 - threadIdx.x + kVecThreads * kCopyVecPerThread <= nvec) {
    u64x4 v[kCopyVecPerThread];
#pragma unroll
    for (int u = 0; u < kCopyVecPerThread; ++u) v[u] = ld256_stream(src + g0 + u * kVecThreads);
#pragma unroll
    for (int u = 0; u < kCopyVecPerThread; ++u) st256_stream(dst + g0 + u * kVecThreads, v[u]);
  } else {
    for (int u = 0; u < kCopyVecPerThread; ++u) {
      const uint64_t g = g0 + (uint64_t)u * kVecThreads;
      if (g < nvec) st256_stream(dst + g, ld256_stream(src + g));
    }
  }
}

// --------------------------------------------------------------- VEC triad --
struct alignas(32) f32x8 {
  float v[8];
};
__device__ __forceinline__ f32x8 ld256_f(const f32x8* p) {
  f32x8 r;
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]),
                 "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st256_f(f32x8* p, const f32x8& r) {
  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
               "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]),
               "f"(r.v[6]), "f"(r.v[7])
               : "memory");
}
__device__ __forceinline__ f32x8 triad8(const f32x8& b, const f32x8& c, float s) {
  f32x8 a;
#pragma unroll
  for (int j = 0; j < 8; ++j) a.v[j] = fmaf(s, c.v[j], b.v[j]);  // single rounding, as the oracle
  return a;
}

__global__ void __launch_bounds__(kVecThreads)
triad_vec_kernel(f32x8* __restrict__ a, const f32x8* __restrict__ b, const f32x8* __restrict__ c,
                 float s, uint64_t nvec) {
  const uint64_t g0 = (uint64_t)blockIdx.x * (kVecThreads * kTriVecPerThread) + threadIdx.x;
  if (g0 - threadIdx.x + kVecThreads * kTriVecPerThread <= nvec) {
    f32x8 bv[kTriVecPerThread], cv[kTriVecPerThread];
#pragma unroll
    for (int u = 0; u < kTriVecPerThread; ++u) {
      bv[u] = ld256_f(b + g0 + u * kVecThreads);
      cv[u] = ld256_f(c + g0 + u * kVecThreads);
    }
#pragma unroll
    for (int u = 0; u < kTriVecPerThread; ++u) st256_f(a + g0 + u * kVecThreads, triad8(bv[u], cv[u], s));
  } else {
    for (int u = 0; u < kTriVecPerThread; ++u) {
      const uint64_t g = g0 + (uint64_t)u * kVecThreads;
      if (g < nvec) st256_f(a + g, triad8(ld256_f(b + g), ld256_f(c + g), s));
    }
  }
}

// ---------------------------------------------------------------- TMA copy --
// One CTA moves kCopyTiles tiles of kTile bytes: one thread arms an mbarrier per tile,
// issues every bulk load (UBLKCP global->shared), then drains each tile shared->global as
// it lands.  No thread touches the data.  64 KiB of shared memory per CTA, three CTAs per SM.
constexpr uint32_t kTile = 16 * 1024;
constexpr int kCopyTiles = 4;
constexpr uint32_t kCopySpan = kTile * kCopyTiles;
constexpr size_t kCopySmem = (size_t)kCopySpan + 128;

__global__ void __launch_bounds__(32)
copy_tma_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[kCopyTiles];
  uint8_t* tiles = reinterpret_cast<uint8_t*>(((uintptr_t)smem + 127) & ~(uintptr_t)127);
  if (threadIdx.x != 0) return;
#pragma unroll
  for (int t = 0; t < kCopyTiles; ++t) mbar_init(&full[t], 1);
  mbar_fence_init();
  const uint64_t base = (uint64_t)blockIdx.x * kCopySpan;
#pragma unroll
  for (int t = 0; t < kCopyTiles; ++t) {
    mbar_expect_tx(&full[t], kTile);
    bulk_g2s(tiles + t * kTile, src + base + (uint64_t)t * kTile, kTile, &full[t]);
  }
#pragma unroll
  for (int t = 0; t < kCopyTiles; ++t) {
    mbar_wait(&full[t], 0);
    bulk_s2g(dst + base + (uint64_t)t * kTile, tiles + t * kTile, kTile);
  }
  bulk_commit();
  bulk_wait_read<0>();  // shared memory must stay valid until the stores have read it
}

// --------------------------------------------------------------- TMA triad --
// One CTA per 16 KiB tile of each array: thread 0 bulk-loads the b and c tiles onto one
// mbarrier; four warps combine them in place in shared memory (LDS.128 / STS.128), fence the
// generic proxy against the async proxy, and thread 0 bulk-stores the result tile.
constexpr uint32_t kTriTile = 16 * 1024;
constexpr int kTriThreads = 128;
constexpr size_t kTriSmem = (size_t)2 * kTriTile + 128;

__global__ void __launch_bounds__(kTriThreads)
triad_tma_kernel(uint8_t* __restrict__ a, const uint8_t* __restrict__ b,
                 const uint8_t* __restrict__ c, float s) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full;
  uint8_t* tiles = reinterpret_cast<uint8_t*>(((uintptr_t)smem + 127) & ~(uintptr_t)127);
  const uint64_t off = (uint64_t)blockIdx.x * kTriTile;
  if (threadIdx.x == 0) {
    mbar_init(&full, 1);
    mbar_fence_init();
    mbar_expect_tx(&full, 2 * kTriTile);
    bulk_g2s(tiles, b + off, kTriTile, &full);
    bulk_g2s(tiles + kTriTile, c + off, kTriTile, &full);
  }
  __syncthreads();  // barrier init visible to the waiters
  mbar_wait(&full, 0);
  float4* sb = reinterpret_cast<float4*>(tiles);
  const float4* sc = reinterpret_cast<const float4*>(tiles + kTriTile);
#pragma unroll
  for (int k = 0; k < (int)(kTriTile / 16 / kTriThreads); ++k) {
    const int idx = threadIdx.x + k * kTriThreads;
    float4 x = sb[idx];
    const float4 y = sc[idx];
    x.x = fmaf(s, y.x, x.x); x.y = fmaf(s, y.y, x.y);
    x.z = fmaf(s, y.z, x.z); x.w = fmaf(s, y.w, x.w);
    sb[idx] = x;
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    bulk_s2g(a + off, tiles, kTriTile);
    bulk_commit();
    bulk_wait_read<0>();
  }
}

// ------------------------------------------------------------- generators ---
__global__ void __launch_bounds__(256) gen_f32_kernel(float* __restrict__ dst, uint64_t n,
                                                      uint64_t seed) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = gen_f32(seed, i);
}
__global__ void __launch_bounds__(256) gen_bf16_kernel(uint16_t* __restrict__ dst, uint64_t n,
                                                       uint64_t seed) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = f32_to_bf16_rne(gen_f32(seed, i));
}
__global__ void __launch_bounds__(256) triad_check_kernel(const float* __restrict__ a, uint64_t n,
                                                          uint64_t seed, float s,
                                                          unsigned long long* __restrict__ count) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned long long bad = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float e = fmaf(s, gen_f32(seed ^ kSeedTriadCXor, i), gen_f32(seed ^ kSeedTriadBXor, i));
    bad += (__float_as_uint(e) != __float_as_uint(a[i]));
  }
  bad = warp_sum_u64(bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(count, bad);
}

// --------------------------------------------------------------- launchers --
cudaError_t stream_kernels_configure() {
  cudaError_t e = cudaFuncSetAttribute(copy_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kCopySmem);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(triad_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)kTriSmem);
}

static unsigned int chunk_grid(uint64_t nvec, uint64_t vec_per_cta) {
  const uint64_t g = (nvec + vec_per_cta - 1) / vec_per_cta;
  return (unsigned int)(g < 1 ? 1 : g);
}

cudaError_t launch_copy(void* dst, const void* src, uint64_t bytes, int variant, int sm_count,
                        cudaStream_t stream) {
  // bytes is a multiple of 32.  The TMA variant moves whole 64 KiB spans; a ragged tail (only
  // the host-buffer parity path has one) goes through the vector kernel.
  (void)sm_count;
  uint64_t done = 0;
  if (variant == GP_STREAM_TMA) {
    const uint64_t nspans = bytes / kCopySpan;
    if (nspans)
      copy_tma_kernel<<<(unsigned int)nspans, 32, kCopySmem, stream>>>(
          reinterpret_cast<const uint8_t*>(src), reinterpret_cast<uint8_t*>(dst));
    done = nspans * kCopySpan;
  }
  if (done < bytes) {
    const uint64_t nvec = (bytes - done) / 32;
    copy_vec_kernel<<<chunk_grid(nvec, kVecThreads * kCopyVecPerThread), kVecThreads, 0, stream>>>(
        reinterpret_cast<const u64x4*>(reinterpret_cast<const uint8_t*>(src) + done),
        reinterpret_cast<u64x4*>(reinterpret_cast<uint8_t*>(dst) + done), nvec);
  }
  return cudaGetLastError();
}

cudaError_t launch_triad(float* a, const float* b, const float* c, float s, uint64_t bytes,
                         int variant, int sm_count, cudaStream_t stream) {
  (void)sm_count;
  uint64_t done = 0;
  uint8_t* a8 = reinterpret_cast<uint8_t*>(a);
  const uint8_t* b8 = reinterpret_cast<const uint8_t*>(b);
  const uint8_t* c8 = reinterpret_cast<const uint8_t*>(c);
  if (variant == GP_STREAM_TMA) {
    const uint64_t ntiles = bytes / kTriTile;
    if (ntiles)
      triad_tma_kernel<<<(unsigned int)ntiles, kTriThreads, kTriSmem, stream>>>(a8, b8, c8, s);
    done = ntiles * kTriTile;
  }
  if (done < bytes) {
    const uint64_t nvec = (bytes - done) / 32;
    triad_vec_kernel<<<chunk_grid(nvec, kVecThreads * kTriVecPerThread), kVecThreads, 0, stream>>>(
        reinterpret_cast<f32x8*>(a8 + done), reinterpret_cast<const f32x8*>(b8 + done),
        reinterpret_cast<const f32x8*>(c8 + done), s, nvec);
  }
  return cudaGetLastError();
}

cudaError_t launch_gen_f32(float* dst, uint64_t n, uint64_t seed, int sm_count, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  gen_f32_kernel<<<sm_count * 8, 256, 0, stream>>>(dst, n, seed);
  return cudaGetLastError();
}
cudaError_t launch_gen_bf16(uint16_t* dst, uint64_t n, uint64_t seed, int sm_count,
                            cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  gen_bf16_kernel<<<sm_count * 8, 256, 0, stream>>>(dst, n, seed);
  return cudaGetLastError();
}
cudaError_t launch_triad_check(const float* a, uint64_t n, uint64_t seed, float s,
                               unsigned long long* d_count, int sm_count, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  triad_check_kernel<<<sm_count * 8, 256, 0, stream>>>(a, n, seed, s, d_count);
  return cudaGetLastError();
}

}  // namespace gp
