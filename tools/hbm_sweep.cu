// hbm_sweep.cu — development tool (not shipped in libgpuprobe.so): sweeps launch shapes and
// memory-instruction variants of the fill / read / copy kernels on one B200 and prints a
// table, so the choices in csrc/pattern.cu and csrc/stream.cu are backed by measurements.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/hbm_sweep tools/hbm_sweep.cu
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <vector>

#include "../k8s-gpu-device-plugin_b200/csrc/common.cuh"
using namespace gp;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

// ---------------------------------------------------------------- stores ----
enum { ST_NA = 0, ST_DEFAULT = 1, ST_CS = 2, ST_EVICT_FIRST = 3, ST_WT = 4 };
template <int MODE>
__device__ __forceinline__ void st256(void* p, const u64x4& r, uint64_t pol) {
  if constexpr (MODE == ST_NA)
    asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(r.v[0]), "l"(r.v[1]), "l"(r.v[2]), "l"(r.v[3]) : "memory");
  else if constexpr (MODE == ST_DEFAULT)
    asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(r.v[0]), "l"(r.v[1]), "l"(r.v[2]), "l"(r.v[3]) : "memory");
  else if constexpr (MODE == ST_CS) {
    asm volatile("st.global.cs.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(r.v[0]), "l"(r.v[1]) : "memory");
    asm volatile("st.global.cs.v2.u64 [%0], {%1,%2};" ::"l"((char*)p + 16), "l"(r.v[2]), "l"(r.v[3]) : "memory");
  } else if constexpr (MODE == ST_EVICT_FIRST)
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.u64 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "l"(r.v[0]), "l"(r.v[1]), "l"(r.v[2]), "l"(r.v[3]), "l"(pol) : "memory");
  else {
    asm volatile("st.global.wt.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(r.v[0]), "l"(r.v[1]) : "memory");
    asm volatile("st.global.wt.v2.u64 [%0], {%1,%2};" ::"l"((char*)p + 16), "l"(r.v[2]), "l"(r.v[3]) : "memory");
  }
}
enum { LD_NA = 0, LD_DEFAULT = 1, LD_EVICT_FIRST = 2, LD_NC = 3 };
template <int MODE>
__device__ __forceinline__ u64x4 ld256(const void* p, uint64_t pol) {
  u64x4 r;
  if constexpr (MODE == LD_NA)
    asm volatile("ld.global.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.v[0]), "=l"(r.v[1]), "=l"(r.v[2]), "=l"(r.v[3]) : "l"(p));
  else if constexpr (MODE == LD_DEFAULT)
    asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.v[0]), "=l"(r.v[1]), "=l"(r.v[2]), "=l"(r.v[3]) : "l"(p));
  else if constexpr (MODE == LD_EVICT_FIRST)
    asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.u64 {%0,%1,%2,%3}, [%4], %5;" : "=l"(r.v[0]), "=l"(r.v[1]), "=l"(r.v[2]), "=l"(r.v[3]) : "l"(p), "l"(pol));
  else
    asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.v[0]), "=l"(r.v[1]), "=l"(r.v[2]), "=l"(r.v[3]) : "l"(p));
  return r;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// grid-stride fill, U stores per loop trip
template <int MODE, int U>
__global__ void fill_gs(uint64_t* w, uint64_t nvec, int pass) {
  const uint64_t pol = policy_evict_first();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; g + (U - 1) * stride < nvec; g += U * stride) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t i = (g + u * stride) << 2;
      u64x4 v;
#pragma unroll
      for (int j = 0; j < 4; ++j) v.v[j] = 1ull << ((unsigned)(i + j + pass) & 63u);
      st256<MODE>(w + i, v, pol);
    }
  }
}
// each CTA owns contiguous chunks of CH bytes, handed out round-robin
template <int MODE>
__global__ void fill_chunk(uint64_t* w, uint64_t nbytes, uint64_t chunk, int pass) {
  const uint64_t pol = policy_evict_first();
  const uint64_t nchunks = nbytes / chunk;
  const uint64_t vec_per_chunk = chunk / 32;
  for (uint64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    for (uint64_t g = threadIdx.x; g < vec_per_chunk; g += blockDim.x) {
      const uint64_t i = (c * vec_per_chunk + g) << 2;
      u64x4 v;
#pragma unroll
      for (int j = 0; j < 4; ++j) v.v[j] = 1ull << ((unsigned)(i + j + pass) & 63u);
      st256<MODE>(w + i, v, pol);
    }
  }
}
// TMA bulk store of a pattern tile kept in shared memory (walking-ones has period 512 B)
template <int TILE>
__global__ void fill_tma(uint8_t* dst, uint64_t ntiles, int pass, int inflight) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* t = reinterpret_cast<uint64_t*>(smem);
  for (int i = threadIdx.x; i < TILE / 8; i += blockDim.x) t[i] = 1ull << ((unsigned)(i + pass) & 63u);
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x != 0) return;
  int pending = 0;
  for (uint64_t n = blockIdx.x; n < ntiles; n += gridDim.x) {
    bulk_s2g(dst + n * TILE, smem, TILE);
    bulk_commit();
    if (++pending >= inflight) { bulk_wait_read<0>(); pending = 0; }
  }
  bulk_wait_all<0>();
}

// read + sum
template <int MODE, int U>
__global__ void read_gs(const uint64_t* w, uint64_t nvec, uint64_t* out) {
  const uint64_t pol = policy_evict_first();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t s = 0, x = 0;
  for (; g + (U - 1) * stride < nvec; g += U * stride) {
    u64x4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld256<MODE>(w + ((g + u * stride) << 2), pol);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      s += v[u].v[0] + v[u].v[1] + v[u].v[2] + v[u].v[3];
      x ^= v[u].v[0] ^ v[u].v[1] ^ v[u].v[2] ^ v[u].v[3];
    }
  }
  s = warp_sum_u64(s) + warp_xor_u64(x);
  if ((threadIdx.x & 31) == 0 && s == 0x1234567) atomicAdd((unsigned long long*)out, (unsigned long long)s);
}
// TMA bulk load into smem ring, consumer warps sum from smem
template <int TILE, int STAGES>
__global__ void read_tma(const uint8_t* src, uint64_t ntiles, uint64_t* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[STAGES], empty[STAGES];
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], blockDim.x - 32); }
    mbar_fence_init();
  }
  __syncthreads();
  const uint64_t first = blockIdx.x, step = gridDim.x;
  const uint64_t mine = first < ntiles ? (ntiles - first + step - 1) / step : 0;
  if (threadIdx.x < 32) {
    if (threadIdx.x != 0) return;
    for (uint64_t n = 0; n < mine; ++n) {
      const int st = (int)(n % STAGES);
      if (n >= STAGES) mbar_wait(&empty[st], (uint32_t)(((n / STAGES) - 1) & 1));
      mbar_expect_tx(&full[st], TILE);
      bulk_g2s(smem + (size_t)st * TILE, src + (first + n * step) * TILE, TILE, &full[st]);
    }
    return;
  }
  const int ct = threadIdx.x - 32, nc = blockDim.x - 32;
  uint64_t s = 0, x = 0;
  for (uint64_t n = 0; n < mine; ++n) {
    const int st = (int)(n % STAGES);
    mbar_wait(&full[st], (uint32_t)((n / STAGES) & 1));
    const ulonglong2* t = reinterpret_cast<const ulonglong2*>(smem + (size_t)st * TILE);
    for (int i = ct; i < TILE / 16; i += nc) { ulonglong2 v = t[i]; s += v.x + v.y; x ^= v.x ^ v.y; }
    mbar_arrive(&empty[st]);
  }
  s = warp_sum_u64(s) + warp_xor_u64(x);
  if ((threadIdx.x & 31) == 0 && s == 0x1234567) atomicAdd((unsigned long long*)out, (unsigned long long)s);
}

// copy
template <int LMODE, int SMODE, int U>
__global__ void copy_gs(const u64x4* src, u64x4* dst, uint64_t nvec) {
  const uint64_t pol = policy_evict_first();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; g + (U - 1) * stride < nvec; g += U * stride) {
    u64x4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld256<LMODE>(src + g + u * stride, pol);
#pragma unroll
    for (int u = 0; u < U; ++u) st256<SMODE>(dst + g + u * stride, v[u], pol);
  }
}
template <int TILE, int STAGES>
__global__ void copy_tma(const uint8_t* src, uint8_t* dst, uint64_t ntiles, int chunked) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[STAGES];
  if (threadIdx.x != 0) return;
  for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
  mbar_fence_init();
  uint64_t first, step, mine;
  if (chunked) {  // contiguous run of tiles per CTA
    const uint64_t per = (ntiles + gridDim.x - 1) / gridDim.x;
    first = blockIdx.x * per; step = 1;
    mine = first < ntiles ? (per < ntiles - first ? per : ntiles - first) : 0;
  } else {
    first = blockIdx.x; step = gridDim.x;
    mine = first < ntiles ? (ntiles - first + step - 1) / step : 0;
  }
  uint64_t issued = 0;
  for (; issued < mine && issued < (uint64_t)(STAGES - 1); ++issued) {
    const int s = (int)(issued % STAGES);
    mbar_expect_tx(&full[s], TILE);
    bulk_g2s(smem + (size_t)s * TILE, src + (first + issued * step) * TILE, TILE, &full[s]);
  }
  for (uint64_t n = 0; n < mine; ++n) {
    const int s = (int)(n % STAGES);
    mbar_wait(&full[s], (uint32_t)((n / STAGES) & 1));
    bulk_s2g(dst + (first + n * step) * TILE, smem + (size_t)s * TILE, TILE);
    bulk_commit();
    if (issued < mine) {
      bulk_wait_read<1>();
      const int r = (int)(issued % STAGES);
      mbar_expect_tx(&full[r], TILE);
      bulk_g2s(smem + (size_t)r * TILE, src + (first + issued * step) * TILE, TILE, &full[r]);
      ++issued;
    }
  }
  bulk_wait_all<0>();
}


// non-persistent: one CTA per contiguous chunk, then exit (memset-style moving window)
template <int MODE, int W>  // W = 32 (256-bit) or 16 (128-bit) bytes per store
__global__ void fill_np(uint8_t* w, uint64_t chunk, int pass, int constant) {
  const uint64_t pol = policy_evict_first();
  uint8_t* base = w + (uint64_t)blockIdx.x * chunk;
  for (uint64_t off = (uint64_t)threadIdx.x * W; off < chunk; off += (uint64_t)blockDim.x * W) {
    const uint64_t i = ((uint64_t)blockIdx.x * chunk + off) >> 3;
    if constexpr (W == 32) {
      u64x4 v;
#pragma unroll
      for (int j = 0; j < 4; ++j) v.v[j] = constant ? 0x0707070707070707ull : 1ull << ((unsigned)(i + j + pass) & 63u);
      st256<MODE>(base + off, v, pol);
    } else {
      uint64_t a = constant ? 0x0707070707070707ull : 1ull << ((unsigned)(i + pass) & 63u);
      uint64_t b = constant ? 0x0707070707070707ull : 1ull << ((unsigned)(i + 1 + pass) & 63u);
      asm volatile("st.global.v2.u64 [%0], {%1,%2};" ::"l"(base + off), "l"(a), "l"(b) : "memory");
    }
  }
}
template <int U>
__global__ void read_np(const uint8_t* w, uint64_t chunk, uint64_t* out) {
  const uint8_t* base = w + (uint64_t)blockIdx.x * chunk;
  uint64_t s = 0, x = 0;
  const uint64_t step = (uint64_t)blockDim.x * 32;
  uint64_t off = (uint64_t)threadIdx.x * 32;
  for (; off + (U - 1) * step < chunk; off += U * step) {
    u64x4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld256<LD_NA>(base + off + u * step, 0);
#pragma unroll
    for (int u = 0; u < U; ++u) { s += v[u].v[0] + v[u].v[1] + v[u].v[2] + v[u].v[3]; x ^= v[u].v[0] ^ v[u].v[1] ^ v[u].v[2] ^ v[u].v[3]; }
  }
  s = warp_sum_u64(s) + warp_xor_u64(x);
  if ((threadIdx.x & 31) == 0 && s == 0x1234567) atomicAdd((unsigned long long*)out, (unsigned long long)s);
}
template <int U>
__global__ void copy_np(const uint8_t* src, uint8_t* dst, uint64_t chunk) {
  const uint64_t b0 = (uint64_t)blockIdx.x * chunk;
  const uint64_t step = (uint64_t)blockDim.x * 32;
  uint64_t off = (uint64_t)threadIdx.x * 32;
  for (; off + (U - 1) * step < chunk; off += U * step) {
    u64x4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld256<LD_NA>(src + b0 + off + u * step, 0);
#pragma unroll
    for (int u = 0; u < U; ++u) st256<ST_NA>(dst + b0 + off + u * step, v[u], 0);
  }
}
// TMA copy with separate load and store threads (two warps), STAGES-deep ring
template <int TILE, int STAGES, int SLAG>
__global__ void copy_tma2(const uint8_t* src, uint8_t* dst, uint64_t ntiles) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[STAGES], empty[STAGES];
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_fence_init();
  }
  __syncthreads();
  const uint64_t first = blockIdx.x, step = gridDim.x;
  const uint64_t mine = first < ntiles ? (ntiles - first + step - 1) / step : 0;
  if (threadIdx.x == 0) {
    for (uint64_t n = 0; n < mine; ++n) {
      const int st = (int)(n % STAGES);
      if (n >= STAGES) mbar_wait(&empty[st], (uint32_t)(((n / STAGES) - 1) & 1));
      mbar_expect_tx(&full[st], TILE);
      bulk_g2s(smem + (size_t)st * TILE, src + (first + n * step) * TILE, TILE, &full[st]);
    }
  } else if (threadIdx.x == 32) {
    for (uint64_t n = 0; n < mine; ++n) {
      const int st = (int)(n % STAGES);
      mbar_wait(&full[st], (uint32_t)((n / STAGES) & 1));
      bulk_s2g(dst + (first + n * step) * TILE, smem + (size_t)st * TILE, TILE);
      bulk_commit();
      if (n >= SLAG) { bulk_wait_read<SLAG>(); mbar_arrive(&empty[(int)((n - SLAG) % STAGES)]); }
    }
    bulk_wait_all<0>();
  }
}

// non-persistent TMA copy: one CTA moves NT tiles of TILE bytes, loads first, then stores
template <int TILE, int NT>
__global__ void copy_tma_np(const uint8_t* src, uint8_t* dst) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[NT];
  if (threadIdx.x != 0) return;
  for (int t = 0; t < NT; ++t) mbar_init(&full[t], 1);
  mbar_fence_init();
  const uint64_t base = (uint64_t)blockIdx.x * TILE * NT;
  for (int t = 0; t < NT; ++t) { mbar_expect_tx(&full[t], TILE); bulk_g2s(smem + t * TILE, src + base + (uint64_t)t * TILE, TILE, &full[t]); }
  for (int t = 0; t < NT; ++t) { mbar_wait(&full[t], 0); bulk_s2g(dst + base + (uint64_t)t * TILE, smem + t * TILE, TILE); }
  bulk_commit();
  bulk_wait_read<0>();
}
// ---- round 2 experiments: can the 1R:1W mix be un-mixed? ----------------------------------
// (a) copy_tma_np with an L2 prefetch of the chunk that the CTA `DIST` places ahead will load
__device__ __forceinline__ void bulk_prefetch_l2(const void* g, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(g), "r"(bytes) : "memory");
}
template <int TILE, int NT, int DIST>
__global__ void copy_tma_np_pf(const uint8_t* src, uint8_t* dst, uint64_t total) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[NT];
  if (threadIdx.x != 0) return;
  for (int t = 0; t < NT; ++t) mbar_init(&full[t], 1);
  mbar_fence_init();
  const uint64_t span = (uint64_t)TILE * NT;
  const uint64_t base = (uint64_t)blockIdx.x * span;
  const uint64_t ahead = base + (uint64_t)DIST * span;
  if (ahead + span <= total)
    for (int t = 0; t < NT; ++t) bulk_prefetch_l2(src + ahead + (uint64_t)t * TILE, TILE);
  for (int t = 0; t < NT; ++t) { mbar_expect_tx(&full[t], TILE); bulk_g2s(smem + t * TILE, src + base + (uint64_t)t * TILE, TILE, &full[t]); }
  for (int t = 0; t < NT; ++t) { mbar_wait(&full[t], 0); bulk_s2g(dst + base + (uint64_t)t * TILE, smem + t * TILE, TILE); }
  bulk_commit();
  bulk_wait_read<0>();
}
// (b) phase separation at L2-window grain: one tiny kernel pulls window w+1 of the source into L2 (reads only),
// the next kernel copies window w (its loads hit L2, DRAM sees writes only)
__global__ void l2_pull_window(const uint8_t* src, uint64_t bytes) {
  const uint64_t off = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16384ull;
  if (off < bytes) bulk_prefetch_l2(src + off, 16384);
}

template <int U>
__global__ void triad_np(float* a, const float* b, const float* c, float s) {
  const uint64_t g0 = ((uint64_t)blockIdx.x * blockDim.x * U + threadIdx.x) * 8;
  float4 bv[U][2], cv[U][2];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint64_t g = g0 + (uint64_t)u * blockDim.x * 8;
    asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(bv[u][0].x), "=f"(bv[u][0].y), "=f"(bv[u][0].z), "=f"(bv[u][0].w), "=f"(bv[u][1].x), "=f"(bv[u][1].y), "=f"(bv[u][1].z), "=f"(bv[u][1].w) : "l"(b + g));
    asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(cv[u][0].x), "=f"(cv[u][0].y), "=f"(cv[u][0].z), "=f"(cv[u][0].w), "=f"(cv[u][1].x), "=f"(cv[u][1].y), "=f"(cv[u][1].z), "=f"(cv[u][1].w) : "l"(c + g));
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint64_t g = g0 + (uint64_t)u * blockDim.x * 8;
    asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(a + g), "f"(fmaf(s, cv[u][0].x, bv[u][0].x)), "f"(fmaf(s, cv[u][0].y, bv[u][0].y)), "f"(fmaf(s, cv[u][0].z, bv[u][0].z)), "f"(fmaf(s, cv[u][0].w, bv[u][0].w)), "f"(fmaf(s, cv[u][1].x, bv[u][1].x)), "f"(fmaf(s, cv[u][1].y, bv[u][1].y)), "f"(fmaf(s, cv[u][1].z, bv[u][1].z)), "f"(fmaf(s, cv[u][1].w, bv[u][1].w)) : "memory");
  }
}
// non-persistent TMA triad: CTA loads one b tile and one c tile, combines in place, bulk-stores
template <int TILE>
__global__ void triad_tma_np(uint8_t* a, const uint8_t* b, const uint8_t* c, float s) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full;
  const uint64_t off = (uint64_t)blockIdx.x * TILE;
  if (threadIdx.x == 0) {
    mbar_init(&full, 1);
    mbar_fence_init();
    mbar_expect_tx(&full, 2 * TILE);
    bulk_g2s(smem, b + off, TILE, &full);
    bulk_g2s(smem + TILE, c + off, TILE, &full);
  }
  __syncthreads();
  mbar_wait(&full, 0);
  float4* sb = reinterpret_cast<float4*>(smem);
  const float4* sc = reinterpret_cast<const float4*>(smem + TILE);
  for (int i = threadIdx.x; i < TILE / 16; i += blockDim.x) {
    float4 x = sb[i]; const float4 y = sc[i];
    x.x = fmaf(s, y.x, x.x); x.y = fmaf(s, y.y, x.y); x.z = fmaf(s, y.z, x.z); x.w = fmaf(s, y.w, x.w);
    sb[i] = x;
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) { bulk_s2g(a + off, smem, TILE); bulk_commit(); bulk_wait_read<0>(); }
}

struct Case {
  std::string name;
  double bytes;
  std::function<void(cudaStream_t)> run;
};

int main(int argc, char** argv) {
  const uint64_t N = (argc > 1 ? strtoull(argv[1], 0, 10) : 8ull) << 30;
  const char* filter = argc > 2 ? argv[2] : "";
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  const int sm = p.multiProcessorCount;
  printf("# %s, %d SMs, N = %llu GiB per array\n", p.name, sm, (unsigned long long)(N >> 30));
  uint8_t *a, *b;
  const uint64_t kPad = 256ull << 20;  // room for the offset experiments (r2 off*)
  CK(cudaMalloc(&a, N)); CK(cudaMalloc(&b, N + kPad));
  CK(cudaMemset(a, 1, N)); CK(cudaMemset(b, 2, N));
  uint64_t* out; CK(cudaMalloc(&out, 8)); CK(cudaMemset(out, 0, 8));
  cudaStream_t s; CK(cudaStreamCreate(&s));
  const uint64_t nvec = N / 32;
  std::vector<Case> cases;
  uint64_t* A = (uint64_t*)a;

#define FILL_GS(MODE, U, BPS, T) cases.push_back({"fill_gs " #MODE " U" #U " bps" #BPS " t" #T, (double)N, [=](cudaStream_t st) { fill_gs<MODE, U><<<sm * BPS, T, 0, st>>>(A, nvec, 3); }})
  FILL_GS(ST_NA, 1, 8, 256); FILL_GS(ST_NA, 1, 4, 512); FILL_GS(ST_NA, 1, 2, 1024); FILL_GS(ST_NA, 1, 4, 256); FILL_GS(ST_NA, 1, 2, 256); FILL_GS(ST_NA, 1, 1, 256);
  FILL_GS(ST_NA, 4, 8, 256); FILL_GS(ST_NA, 4, 2, 256); FILL_GS(ST_NA, 1, 16, 128); FILL_GS(ST_NA, 1, 1, 1024);
  FILL_GS(ST_DEFAULT, 1, 8, 256); FILL_GS(ST_CS, 1, 8, 256); FILL_GS(ST_EVICT_FIRST, 1, 8, 256); FILL_GS(ST_WT, 1, 8, 256);
  FILL_GS(ST_EVICT_FIRST, 1, 2, 256); FILL_GS(ST_DEFAULT, 1, 2, 256);
#define FILL_CH(MODE, CH, BPS, T) cases.push_back({"fill_chunk " #MODE " ch" #CH " bps" #BPS " t" #T, (double)N, [=](cudaStream_t st) { fill_chunk<MODE><<<sm * BPS, T, 0, st>>>(A, N, CH, 3); }})
  FILL_CH(ST_NA, 65536, 4, 256); FILL_CH(ST_NA, 1048576, 4, 256); FILL_CH(ST_NA, 16384, 8, 256); FILL_CH(ST_NA, 262144, 2, 512);
#define FILL_TMA(TILE, BPS, INF) { CK(cudaFuncSetAttribute(fill_tma<TILE>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE)); cases.push_back({"fill_tma tile" #TILE " bps" #BPS " inflight" #INF, (double)N, [=](cudaStream_t st) { fill_tma<TILE><<<sm * BPS, 128, TILE, st>>>(a, N / TILE, 3, INF); }}); }
  FILL_TMA(32768, 1, 4); FILL_TMA(32768, 1, 16); FILL_TMA(32768, 2, 8); FILL_TMA(65536, 1, 8); FILL_TMA(16384, 4, 8); FILL_TMA(8192, 8, 8); FILL_TMA(65536, 2, 4);

#define READ_GS(MODE, U, BPS, T) cases.push_back({"read_gs " #MODE " U" #U " bps" #BPS " t" #T, (double)N, [=](cudaStream_t st) { read_gs<MODE, U><<<sm * BPS, T, 0, st>>>(A, nvec, out); }})
  READ_GS(LD_NA, 4, 4, 256); READ_GS(LD_NA, 4, 8, 256); READ_GS(LD_NA, 8, 4, 256); READ_GS(LD_NA, 2, 8, 256); READ_GS(LD_NA, 8, 2, 256); READ_GS(LD_NA, 4, 2, 512); READ_GS(LD_NA, 4, 2, 1024);
  READ_GS(LD_DEFAULT, 4, 4, 256); READ_GS(LD_EVICT_FIRST, 4, 4, 256); READ_GS(LD_NC, 4, 4, 256); READ_GS(LD_NA, 1, 8, 256); READ_GS(LD_NA, 8, 8, 256);
#define READ_TMA(TILE, STAGES, T) { CK(cudaFuncSetAttribute(read_tma<TILE, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE * STAGES)); cases.push_back({"read_tma tile" #TILE " st" #STAGES " t" #T, (double)N, [=](cudaStream_t st) { read_tma<TILE, STAGES><<<sm, T, TILE * STAGES, st>>>(a, N / TILE, out); }}); }
  READ_TMA(32768, 6, 160); READ_TMA(16384, 12, 160); READ_TMA(32768, 6, 288); READ_TMA(65536, 3, 288);

#define COPY_GS(L, S, U, BPS, T) cases.push_back({"copy_gs " #L " " #S " U" #U " bps" #BPS " t" #T, 2.0 * N, [=](cudaStream_t st) { copy_gs<L, S, U><<<sm * BPS, T, 0, st>>>((const u64x4*)a, (u64x4*)b, nvec); }})
  COPY_GS(LD_NA, ST_NA, 4, 4, 256); COPY_GS(LD_NA, ST_NA, 4, 2, 256); COPY_GS(LD_NA, ST_NA, 8, 2, 256); COPY_GS(LD_NA, ST_NA, 2, 8, 256); COPY_GS(LD_NA, ST_NA, 1, 8, 256); COPY_GS(LD_NA, ST_NA, 4, 1, 1024); COPY_GS(LD_NA, ST_NA, 8, 4, 256);
  COPY_GS(LD_EVICT_FIRST, ST_EVICT_FIRST, 4, 4, 256); COPY_GS(LD_DEFAULT, ST_DEFAULT, 4, 4, 256); COPY_GS(LD_NA, ST_CS, 4, 4, 256); COPY_GS(LD_NC, ST_NA, 4, 4, 256); COPY_GS(LD_EVICT_FIRST, ST_NA, 4, 4, 256);
#define COPY_TMA(TILE, STAGES, BPS, CH) { CK(cudaFuncSetAttribute(copy_tma<TILE, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE * STAGES)); cases.push_back({"copy_tma tile" #TILE " st" #STAGES " bps" #BPS " chunked" #CH, 2.0 * N, [=](cudaStream_t st) { copy_tma<TILE, STAGES><<<sm * BPS, 32, TILE * STAGES, st>>>(a, b, N / TILE, CH); }}); }
  COPY_TMA(32768, 6, 1, 0); COPY_TMA(16384, 12, 1, 0); COPY_TMA(65536, 3, 1, 0); COPY_TMA(8192, 24, 1, 0); COPY_TMA(32768, 3, 2, 0); COPY_TMA(16384, 6, 2, 0); COPY_TMA(16384, 3, 4, 0); COPY_TMA(8192, 3, 8, 0);
  COPY_TMA(32768, 6, 1, 1); COPY_TMA(32768, 4, 1, 0); COPY_TMA(32768, 2, 1, 0); COPY_TMA(16384, 4, 1, 0);

#define FILL_NP(MODE, W, CH, T, C) cases.push_back({"fill_np " #MODE " w" #W " ch" #CH " t" #T " const" #C, (double)N, [=](cudaStream_t st) { fill_np<MODE, W><<<(unsigned)(N / CH), T, 0, st>>>(a, CH, 3, C); }})
  FILL_NP(ST_NA, 32, 8192, 256, 0); FILL_NP(ST_NA, 32, 16384, 256, 0); FILL_NP(ST_NA, 32, 32768, 256, 0); FILL_NP(ST_NA, 32, 65536, 512, 0); FILL_NP(ST_NA, 32, 16384, 512, 0); FILL_NP(ST_NA, 32, 32768, 1024, 0); FILL_NP(ST_NA, 32, 4096, 128, 0);
  FILL_NP(ST_NA, 32, 16384, 256, 1); FILL_NP(ST_DEFAULT, 32, 16384, 256, 0); FILL_NP(ST_DEFAULT, 16, 16384, 256, 0); FILL_NP(ST_DEFAULT, 16, 16384, 256, 1); FILL_NP(ST_DEFAULT, 16, 8192, 512, 0); FILL_NP(ST_DEFAULT, 16, 4096, 256, 0); FILL_NP(ST_DEFAULT, 16, 65536, 1024, 0);
  FILL_NP(ST_NA, 32, 1048576, 1024, 0); FILL_NP(ST_NA, 32, 262144, 1024, 0);
#define READ_NP(U, CH, T) cases.push_back({"read_np U" #U " ch" #CH " t" #T, (double)N, [=](cudaStream_t st) { read_np<U><<<(unsigned)(N / CH), T, 0, st>>>(a, CH, out); }})
  READ_NP(4, 32768, 256); READ_NP(4, 65536, 512); READ_NP(2, 16384, 256); READ_NP(4, 131072, 1024); READ_NP(8, 65536, 256); READ_NP(1, 8192, 256); READ_NP(4, 1048576, 1024); READ_NP(2, 65536, 1024);
#define COPY_NP(U, CH, T) cases.push_back({"copy_np U" #U " ch" #CH " t" #T, 2.0 * N, [=](cudaStream_t st) { copy_np<U><<<(unsigned)(N / CH), T, 0, st>>>(a, b, CH); }})
  COPY_NP(4, 32768, 256); COPY_NP(4, 65536, 512); COPY_NP(2, 16384, 256); COPY_NP(4, 131072, 1024); COPY_NP(8, 65536, 256); COPY_NP(1, 8192, 256); COPY_NP(4, 1048576, 1024); COPY_NP(8, 131072, 512);
#define COPY_TMA2(TILE, STAGES, SLAG, BPS) { CK(cudaFuncSetAttribute(copy_tma2<TILE, STAGES, SLAG>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE * STAGES)); cases.push_back({"copy_tma2 tile" #TILE " st" #STAGES " slag" #SLAG " bps" #BPS, 2.0 * N, [=](cudaStream_t st) { copy_tma2<TILE, STAGES, SLAG><<<sm * BPS, 64, TILE * STAGES, st>>>(a, b, N / TILE); }}); }
  COPY_TMA2(16384, 4, 1, 1); COPY_TMA2(16384, 6, 1, 1); COPY_TMA2(16384, 8, 2, 1); COPY_TMA2(16384, 12, 2, 1); COPY_TMA2(32768, 4, 1, 1); COPY_TMA2(32768, 6, 2, 1); COPY_TMA2(8192, 8, 2, 1); COPY_TMA2(8192, 16, 4, 1); COPY_TMA2(16384, 4, 1, 2); COPY_TMA2(8192, 4, 1, 4); COPY_TMA2(65536, 3, 1, 1); COPY_TMA2(4096, 16, 4, 1);
  COPY_TMA(16384, 3, 1, 0); COPY_TMA(16384, 5, 1, 0); COPY_TMA(16384, 6, 1, 0); COPY_TMA(16384, 8, 1, 0); COPY_TMA(8192, 4, 1, 0); COPY_TMA(8192, 6, 1, 0); COPY_TMA(8192, 8, 1, 0); COPY_TMA(8192, 12, 1, 0); COPY_TMA(4096, 8, 1, 0); COPY_TMA(4096, 16, 1, 0); COPY_TMA(16384, 4, 2, 0); COPY_TMA(8192, 4, 2, 0); COPY_TMA(16384, 4, 1, 1); COPY_TMA(32768, 3, 1, 0); COPY_TMA(24576, 4, 1, 0); COPY_TMA(12288, 4, 1, 0); COPY_TMA(12288, 6, 1, 0);
  cases.push_back({"cudaMemsetAsync 0", (double)N, [=](cudaStream_t st) { cudaMemsetAsync(a, 0, N, st); }});

#define COPY_TMA_NP(TILE, NT) { CK(cudaFuncSetAttribute(copy_tma_np<TILE, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE * NT)); cases.push_back({"copy_tma_np tile" #TILE " nt" #NT, 2.0 * N, [=](cudaStream_t st) { copy_tma_np<TILE, NT><<<(unsigned)(N / (TILE * NT)), 32, TILE * NT, st>>>(a, b); }}); }
  COPY_TMA_NP(16384, 1); COPY_TMA_NP(16384, 2); COPY_TMA_NP(16384, 4); COPY_TMA_NP(32768, 1); COPY_TMA_NP(32768, 2); COPY_TMA_NP(8192, 1); COPY_TMA_NP(8192, 2); COPY_TMA_NP(8192, 4); COPY_TMA_NP(65536, 1); COPY_TMA_NP(4096, 4); COPY_TMA_NP(4096, 1); COPY_TMA_NP(24576, 1);
#define COPY_TMA_NP_PF(TILE, NT, DIST) { CK(cudaFuncSetAttribute(copy_tma_np_pf<TILE, NT, DIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE * NT)); cases.push_back({"r2 copy_tma_np+prefetchL2 tile" #TILE " nt" #NT " dist" #DIST, 2.0 * N, [=](cudaStream_t st) { copy_tma_np_pf<TILE, NT, DIST><<<(unsigned)(N / (TILE * NT)), 32, TILE * NT, st>>>(a, b, N); }}); }
  COPY_TMA_NP(16384, 4);  // the shipped shape, as the baseline of this block
  COPY_TMA_NP_PF(16384, 4, 148); COPY_TMA_NP_PF(16384, 4, 512); COPY_TMA_NP_PF(16384, 4, 1024); COPY_TMA_NP_PF(16384, 4, 4096); COPY_TMA_NP_PF(16384, 4, 64);
#define COPY_WINDOWED(WMIB) cases.push_back({"r2 copy windowed: pull " #WMIB " MiB into L2, then copy it", 2.0 * N, [=](cudaStream_t st) { \
    const uint64_t W = (uint64_t)WMIB << 20; \
    l2_pull_window<<<(unsigned)(W / 16384 / 64), 64, 0, st>>>(a, W); \
    for (uint64_t off = 0; off + W <= N; off += W) { /* W must divide N: a partial window is skipped, never overrun */ \
      if (off + 2 * W <= N) l2_pull_window<<<(unsigned)(W / 16384 / 64), 64, 0, st>>>(a + off + W, W); \
      copy_tma_np<16384, 4><<<(unsigned)(W / 65536), 32, 65536, st>>>(a + off, b + off); } }})
  COPY_WINDOWED(32); COPY_WINDOWED(64);
  // (c) does the distance between the arrays matter (a power-of-two distance could put src[i] and dst[i] in the same
  // bank)?  Same kernel, destination shifted by an odd number of KiB / MiB inside a padded allocation.
  printf("# a = %p  b = %p  (b - a = %lld MiB)\n", (void*)a, (void*)b, (long long)((b - a) >> 20));
#define COPY_OFF(OFFKIB) cases.push_back({"r2 off copy_tma_np tile16384 nt4, dst + " #OFFKIB " KiB", 2.0 * N, [=](cudaStream_t st) { copy_tma_np<16384, 4><<<(unsigned)(N / 65536), 32, 65536, st>>>(a, b + (uint64_t)OFFKIB * 1024); }})
  COPY_OFF(0); COPY_OFF(64); COPY_OFF(1088); COPY_OFF(4160); COPY_OFF(33856); COPY_OFF(133184); COPY_OFF(16); COPY_OFF(2048);  // (96 MiB does not divide the array; 128 MiB no longer fits L2)
  { const uint64_t NT3 = N / 2; float* fa = (float*)a; float* fb = (float*)b; float* fc = (float*)(b + NT3);
#define TRIAD_NP(U, T) cases.push_back({"triad_np U" #U " t" #T, 3.0 * NT3, [=](cudaStream_t st) { triad_np<U><<<(unsigned)(NT3 / 32 / (T * U)), T, 0, st>>>(fa, fb, fc, 3.0f); }})
    TRIAD_NP(1, 256); TRIAD_NP(2, 256); TRIAD_NP(4, 256); TRIAD_NP(2, 512); TRIAD_NP(1, 512); TRIAD_NP(2, 128); TRIAD_NP(1, 1024);
#define TRIAD_TMA_NP(TILE, T) { CK(cudaFuncSetAttribute(triad_tma_np<TILE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * TILE)); cases.push_back({"triad_tma_np tile" #TILE " t" #T, 3.0 * NT3, [=](cudaStream_t st) { triad_tma_np<TILE><<<(unsigned)(NT3 / TILE), T, 2 * TILE, st>>>(a, b, b + NT3, 3.0f); }}); }
    TRIAD_TMA_NP(16384, 128); TRIAD_TMA_NP(8192, 128); TRIAD_TMA_NP(32768, 256); TRIAD_TMA_NP(16384, 256); TRIAD_TMA_NP(8192, 64); TRIAD_TMA_NP(4096, 64);
  }
  cases.push_back({"cudaMemcpyAsync D2D", 2.0 * N, [=](cudaStream_t st) { cudaMemcpyAsync(b, a, N, cudaMemcpyDeviceToDevice, st); }});
  cases.push_back({"cudaMemsetAsync", (double)N, [=](cudaStream_t st) { cudaMemsetAsync(a, 7, N, st); }});

  cudaEvent_t ev[12];
  for (auto& e : ev) CK(cudaEventCreate(&e));
  for (auto& c : cases) {
    if (*filter && c.name.find(filter) == std::string::npos) continue;
    for (int i = 0; i < 2; ++i) c.run(s);
    CK(cudaStreamSynchronize(s));
    CK(cudaGetLastError());
    const int iters = 7;
    for (int i = 0; i < iters; ++i) { CK(cudaEventRecord(ev[i], s)); c.run(s); }
    CK(cudaEventRecord(ev[iters], s));
    CK(cudaStreamSynchronize(s));
    CK(cudaGetLastError());
    std::vector<float> ms;
    for (int i = 0; i < iters; ++i) { float t; CK(cudaEventElapsedTime(&t, ev[i], ev[i + 1])); ms.push_back(t); }
    std::sort(ms.begin(), ms.end());
    printf("%-58s best %8.3f ms %8.1f GB/s   median %8.1f GB/s\n", c.name.c_str(), ms[0], c.bytes / ms[0] / 1e6, c.bytes / ms[iters / 2] / 1e6);
    fflush(stdout);
  }
  return 0;
}
