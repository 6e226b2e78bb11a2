// stream.cu — HBM stream probes: copy (K1a) and triad (K1b) of SURVEY.md §8.
//
// No reference counterpart (benchmark/benchmark.go:54-89 profiles the Go
// runtime, it moves no device memory).  Algorithmic bytes per launch
// (SURVEY.md §8d): copy 2N, triad 3N, N = bytes per array.
//
// Two hand-written variants of each, selected by gp_config.stream_variant:
//   VEC : persistent grid, four 256-bit loads in flight per thread, then four
//         256-bit stores (LDG.E.256 / STG.E.256, L1 no-allocate).
//   TMA : one CTA per SM.  A producer thread streams 32 KiB tiles
//         global->shared with cp.async.bulk (UBLKCP) completing on an
//         mbarrier ring; copy drains each tile shared->global with
//         cp.async.bulk from the same thread (no thread ever touches the
//         data); triad lets four consumer warps combine the b and c tiles in
//         place in shared memory (LDS.128/STS.128), fences the generic proxy
//         against the async proxy and bulk-stores the result tile.
#include "common.cuh"
#include "kernels.h"

namespace gp {

constexpr int kVecThreads = 256;

// ---------------------------------------------------------------- VEC copy --
__global__ void __launch_bounds__(kVecThreads)
copy_vec_kernel(const u64x4* __restrict__ src, u64x4* __restrict__ dst, uint64_t nvec) {
  const uint64_t stride = (uint64_t)gridDim.x * kVecThreads;
  uint64_t g = (uint64_t)blockIdx.x * kVecThreads + threadIdx.x;
  for (; g + 3 * stride < nvec; g += 4 * stride) {
    const u64x4 v0 = ld256_stream(src + g);
    const u64x4 v1 = ld256_stream(src + g + stride);
    const u64x4 v2 = ld256_stream(src + g + 2 * stride);
    const u64x4 v3 = ld256_stream(src + g + 3 * stride);
    st256_stream(dst + g, v0);
    st256_stream(dst + g + stride, v1);
    st256_stream(dst + g + 2 * stride, v2);
    st256_stream(dst + g + 3 * stride, v3);
  }
  for (; g < nvec; g += stride) st256_stream(dst + g, ld256_stream(src + g));
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
  const uint64_t stride = (uint64_t)gridDim.x * kVecThreads;
  uint64_t g = (uint64_t)blockIdx.x * kVecThreads + threadIdx.x;
  for (; g + stride < nvec; g += 2 * stride) {
    const f32x8 b0 = ld256_f(b + g), c0 = ld256_f(c + g);
    const f32x8 b1 = ld256_f(b + g + stride), c1 = ld256_f(c + g + stride);
    st256_f(a + g, triad8(b0, c0, s));
    st256_f(a + g + stride, triad8(b1, c1, s));
  }
  for (; g < nvec; g += stride) st256_f(a + g, triad8(ld256_f(b + g), ld256_f(c + g), s));
}

// ---------------------------------------------------------------- TMA copy --
constexpr uint32_t kTile = 32 * 1024;  // bytes per bulk copy
constexpr int kCopyStages = 6;         // 192 KiB of shared memory in flight per SM
constexpr size_t kCopySmem = (size_t)kCopyStages * kTile + 1024;

__global__ void __launch_bounds__(32, 1)
copy_tma_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint64_t ntiles) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[kCopyStages];
  uint8_t* tiles = reinterpret_cast<uint8_t*>(((uintptr_t)smem + 127) & ~(uintptr_t)127);
  if (threadIdx.x != 0) return;
  for (int s = 0; s < kCopyStages; ++s) mbar_init(&full[s], 1);
  mbar_fence_init();

  const uint64_t first = blockIdx.x, step = gridDim.x;
  const uint64_t mine = first < ntiles ? (ntiles - first + step - 1) / step : 0;
  // prologue: kCopyStages-1 loads in flight
  uint64_t issued = 0;
  for (; issued < mine && issued < (uint64_t)(kCopyStages - 1); ++issued) {
    const int s = (int)(issued % kCopyStages);
    mbar_expect_tx(&full[s], kTile);
    bulk_g2s(tiles + (size_t)s * kTile, src + (first + issued * step) * kTile, kTile, &full[s]);
  }
  for (uint64_t n = 0; n < mine; ++n) {
    const int s = (int)(n % kCopyStages);
    mbar_wait(&full[s], (uint32_t)((n / kCopyStages) & 1));
    bulk_s2g(dst + (first + n * step) * kTile, tiles + (size_t)s * kTile, kTile);
    bulk_commit();
    if (issued < mine) {
      // the stage to refill held tile n-1, whose store is the second-youngest group
      bulk_wait_read<1>();
      const int r = (int)(issued % kCopyStages);
      mbar_expect_tx(&full[r], kTile);
      bulk_g2s(tiles + (size_t)r * kTile, src + (first + issued * step) * kTile, kTile, &full[r]);
      ++issued;
    }
  }
  bulk_wait_all<0>();
}

// --------------------------------------------------------------- TMA triad --
constexpr uint32_t kTriTile = 16 * 1024;  // bytes of each of b, c per stage; result overwrites b
constexpr int kTriStages = 6;             // 6 * 32 KiB = 192 KiB
constexpr int kTriConsumers = 128;
constexpr size_t kTriSmem = (size_t)kTriStages * 2 * kTriTile + 1024;

__global__ void __launch_bounds__(32 + kTriConsumers, 1)
triad_tma_kernel(uint8_t* __restrict__ a, const uint8_t* __restrict__ b,
                 const uint8_t* __restrict__ c, float s, uint64_t ntiles) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[kTriStages];
  __shared__ __align__(8) uint64_t empty[kTriStages];
  uint8_t* tiles = reinterpret_cast<uint8_t*>(((uintptr_t)smem + 127) & ~(uintptr_t)127);
  if (threadIdx.x == 0) {
    for (int i = 0; i < kTriStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_fence_init();
  }
  __syncthreads();
  const uint64_t first = blockIdx.x, step = gridDim.x;
  const uint64_t mine = first < ntiles ? (ntiles - first + step - 1) / step : 0;

  if (threadIdx.x < 32) {
    if (threadIdx.x != 0) return;
    for (uint64_t n = 0; n < mine; ++n) {  // producer
      const int st = (int)(n % kTriStages);
      if (n >= (uint64_t)kTriStages) mbar_wait(&empty[st], (uint32_t)(((n / kTriStages) - 1) & 1));
      uint8_t* sb = tiles + (size_t)st * 2 * kTriTile;
      const uint64_t off = (first + n * step) * kTriTile;
      mbar_expect_tx(&full[st], 2 * kTriTile);
      bulk_g2s(sb, b + off, kTriTile, &full[st]);
      bulk_g2s(sb + kTriTile, c + off, kTriTile, &full[st]);
    }
    return;
  }
  const int ct = threadIdx.x - 32;  // consumer thread 0..127
  for (uint64_t n = 0; n < mine; ++n) {
    const int st = (int)(n % kTriStages);
    mbar_wait(&full[st], (uint32_t)((n / kTriStages) & 1));
    float4* sb = reinterpret_cast<float4*>(tiles + (size_t)st * 2 * kTriTile);
    const float4* sc = reinterpret_cast<const float4*>(tiles + (size_t)st * 2 * kTriTile + kTriTile);
#pragma unroll
    for (int k = 0; k < (int)(kTriTile / 16 / kTriConsumers); ++k) {
      const int idx = ct + k * kTriConsumers;
      float4 x = sb[idx];
      const float4 y = sc[idx];
      x.x = fmaf(s, y.x, x.x); x.y = fmaf(s, y.y, x.y);
      x.z = fmaf(s, y.z, x.z); x.w = fmaf(s, y.w, x.w);
      sb[idx] = x;
    }
    fence_proxy_async_smem();
    asm volatile("bar.sync 1, %0;" ::"n"(kTriConsumers) : "memory");
    if (ct == 0) {
      bulk_s2g(a + (first + n * step) * kTriTile, sb, kTriTile);
      bulk_commit();
      if (n > 0) {  // tile n-1's store has finished reading its stage: hand it back
        bulk_wait_read<1>();
        mbar_arrive(&empty[(int)((n - 1) % kTriStages)]);
      }
    }
  }
  if (ct == 0) bulk_wait_all<0>();
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

template <auto kernel>
static int resident_blocks_per_sm() {
  static int cached = 0;
  if (cached == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kVecThreads, 0) != cudaSuccess || n < 1) n = 1;
    cached = n;
  }
  return cached;
}

static int vec_grid(uint64_t nvec, int sm_count, int blocks_per_sm) {
  uint64_t want = (nvec + kVecThreads - 1) / kVecThreads;
  const uint64_t cap = (uint64_t)sm_count * blocks_per_sm;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

cudaError_t launch_copy(void* dst, const void* src, uint64_t bytes, int variant, int sm_count,
                        cudaStream_t stream) {
  // bytes is a multiple of 32.  The TMA variant moves whole 32 KiB tiles; a ragged tail (only
  // the host-buffer parity path has one) goes through the vector kernel.
  uint64_t done = 0;
  if (variant == GP_STREAM_TMA) {
    const uint64_t ntiles = bytes / kTile;
    if (ntiles) {
      const int grid = (int)(ntiles < (uint64_t)sm_count ? ntiles : (uint64_t)sm_count);
      copy_tma_kernel<<<grid, 32, kCopySmem, stream>>>(reinterpret_cast<const uint8_t*>(src),
                                                       reinterpret_cast<uint8_t*>(dst), ntiles);
    }
    done = ntiles * kTile;
  }
  if (done < bytes) {
    const uint64_t nvec = (bytes - done) / 32;
    copy_vec_kernel<<<vec_grid(nvec, sm_count, resident_blocks_per_sm<copy_vec_kernel>()), kVecThreads, 0, stream>>>(
        reinterpret_cast<const u64x4*>(reinterpret_cast<const uint8_t*>(src) + done),
        reinterpret_cast<u64x4*>(reinterpret_cast<uint8_t*>(dst) + done), nvec);
  }
  return cudaGetLastError();
}

cudaError_t launch_triad(float* a, const float* b, const float* c, float s, uint64_t bytes,
                         int variant, int sm_count, cudaStream_t stream) {
  uint64_t done = 0;
  uint8_t* a8 = reinterpret_cast<uint8_t*>(a);
  const uint8_t* b8 = reinterpret_cast<const uint8_t*>(b);
  const uint8_t* c8 = reinterpret_cast<const uint8_t*>(c);
  if (variant == GP_STREAM_TMA) {
    const uint64_t ntiles = bytes / kTriTile;
    if (ntiles) {
      const int grid = (int)(ntiles < (uint64_t)sm_count ? ntiles : (uint64_t)sm_count);
      triad_tma_kernel<<<grid, 32 + kTriConsumers, kTriSmem, stream>>>(a8, b8, c8, s, ntiles);
    }
    done = ntiles * kTriTile;
  }
  if (done < bytes) {
    const uint64_t nvec = (bytes - done) / 32;
    triad_vec_kernel<<<vec_grid(nvec, sm_count, resident_blocks_per_sm<triad_vec_kernel>()), kVecThreads, 0, stream>>>(
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
