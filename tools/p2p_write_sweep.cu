// p2p_write_sweep — why do in-kernel peer WRITES stop at ~713 GB/s when the copy engine reaches ~778?
// GPU 0 writes 1 GiB into GPU 1 with every store shape we can think of; best of 3, CUDA events on GPU 0.
// Development tool; results in profiles/p2p_write_sweep_r02.txt.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/p2p_write_sweep tools/p2p_write_sweep.cu
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstdio>
#include <functional>
#include <string>
#include <vector>

enum Op { WB, CG, CS, WT, NOALLOC, RELAXED_SYS, EVICT_FIRST };

template <int BYTES, int OP>
__device__ __forceinline__ void store(void* p, uint64_t a, uint64_t b, uint64_t c, uint64_t d, uint64_t pol) {
  if constexpr (BYTES == 32) {
    if constexpr (OP == WB) asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
    else if constexpr (OP == NOALLOC) asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
    else if constexpr (OP == EVICT_FIRST) asm volatile("st.global.L2::cache_hint.v4.u64 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d), "l"(pol) : "memory");
    else asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
  } else if constexpr (BYTES == 16) {
    if constexpr (OP == WB) asm volatile("st.global.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
    else if constexpr (OP == CG) asm volatile("st.global.cg.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
    else if constexpr (OP == CS) asm volatile("st.global.cs.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
    else if constexpr (OP == WT) asm volatile("st.global.wt.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
    else if constexpr (OP == NOALLOC) asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
    else if constexpr (OP == RELAXED_SYS) asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
    else asm volatile("st.global.L2::cache_hint.v2.u64 [%0], {%1,%2}, %3;" ::"l"(p), "l"(a), "l"(b), "l"(pol) : "memory");
  } else {
    if constexpr (OP == WB) asm volatile("st.global.u64 [%0], %1;" ::"l"(p), "l"(a) : "memory");
    else if constexpr (OP == RELAXED_SYS) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(a) : "memory");
    else asm volatile("st.global.cg.u64 [%0], %1;" ::"l"(p), "l"(a) : "memory");
  }
}

// grid-stride, every thread stores BYTES per trip; unroll U independent stores per trip
template <int BYTES, int OP, int U>
__global__ void wr_kernel(uint8_t* __restrict__ peer, uint64_t bytes) {
  uint64_t pol = 0;
  if constexpr (OP == EVICT_FIRST) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  const uint64_t n = bytes / BYTES;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; g + (U - 1) * stride < n; g += U * stride) {
#pragma unroll
    for (int u = 0; u < U; ++u) { const uint64_t i = g + u * stride; store<BYTES, OP>(peer + i * BYTES, i, ~i, i ^ 0x55, i + 7, pol); }
  }
  for (; g < n; g += stride) store<BYTES, OP>(peer + g * BYTES, g, ~g, g ^ 0x55, g + 7, pol);
}

// TMA bulk store: TILE bytes per cp.async.bulk, NBUF tiles in flight per CTA
template <int TILE, int NBUF>
__global__ void wr_tma_kernel(uint8_t* __restrict__ peer, uint64_t bytes) {
  extern __shared__ __align__(128) uint8_t sm[];
  for (int i = threadIdx.x; i < TILE * NBUF / 8; i += blockDim.x) reinterpret_cast<uint64_t*>(sm)[i] = i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint64_t ntiles = bytes / TILE;
    int b = 0;
    for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(peer + t * TILE),
                   "r"((uint32_t)__cvta_generic_to_shared(sm + b * TILE)), "r"(TILE) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NBUF - 1) : "memory");
      b = (b + 1) % NBUF;
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

struct Case { std::string name; std::function<void(cudaStream_t)> run; };

int main() {
  int n = 0;
  cudaGetDeviceCount(&n);
  if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
  const uint64_t bytes = 1ull << 30;
  uint8_t *d0, *d1;
  cudaSetDevice(1); cudaMalloc(&d1, bytes); cudaDeviceEnablePeerAccess(0, 0);
  cudaSetDevice(0); cudaMalloc(&d0, bytes); cudaDeviceEnablePeerAccess(1, 0);
  cudaMemset(d0, 1, bytes);
  cudaStream_t s0; cudaStreamCreate(&s0);
  cudaSetDevice(1); cudaStream_t s1; cudaStreamCreate(&s1); cudaSetDevice(0);
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  std::vector<Case> cases;
#define K(B, OP, U, GRID, THR) cases.push_back({std::string("st" #B " " #OP " unroll" #U " grid=" #GRID " thr=" #THR), [=](cudaStream_t s) { wr_kernel<B, OP, U><<<(GRID), (THR), 0, s>>>(d1, bytes); }})
  const int np512 = (int)(bytes / 32 / 512);
  K(32, NOALLOC, 1, np512, 512);   // the library's shape
  K(32, WB, 1, np512, 512);
  K(32, EVICT_FIRST, 1, np512, 512);
  K(32, NOALLOC, 4, sms * 4, 512);
  K(32, NOALLOC, 4, sms * 2, 1024);
  K(32, NOALLOC, 8, sms, 1024);
  K(32, NOALLOC, 1, sms * 16, 128);
  K(16, WB, 1, (int)(bytes / 16 / 512), 512);
  K(16, WB, 4, sms * 4, 512);
  K(16, CG, 4, sms * 4, 512);
  K(16, CS, 4, sms * 4, 512);
  K(16, WT, 4, sms * 4, 512);
  K(16, NOALLOC, 4, sms * 4, 512);
  K(16, RELAXED_SYS, 4, sms * 4, 512);
  K(16, EVICT_FIRST, 4, sms * 4, 512);
  K(16, WB, 8, sms * 2, 1024);
  K(16, WB, 2, sms * 8, 256);
  K(8, WB, 4, sms * 4, 512);
  K(8, WB, 8, sms * 4, 1024);
  K(8, RELAXED_SYS, 8, sms * 4, 512);
  K(8, CG, 8, sms * 4, 512);
#define T(TILE, NBUF, GRID) cases.push_back({std::string("tma tile=" #TILE " inflight=" #NBUF " grid=" #GRID), [=](cudaStream_t s) { \
    cudaFuncSetAttribute(wr_tma_kernel<TILE, NBUF>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE * NBUF); \
    wr_tma_kernel<TILE, NBUF><<<(GRID), 128, TILE * NBUF, s>>>(d1, bytes); }})
  T(4096, 4, sms * 4);
  T(16384, 2, sms * 4);
  T(16384, 4, sms * 2);
  T(32768, 4, sms);
  T(65536, 2, sms);
  T(65536, 3, sms);
  T(16384, 8, sms);
  T(1024, 8, sms * 8);
  T(256, 16, sms * 8);
  cases.push_back({"copy engine, src stream (push)", [=](cudaStream_t s) { cudaMemcpyPeerAsync(d1, 1, d0, 0, bytes, s); }});
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (auto& c : cases) {
    float best = 1e30f;
    for (int it = 0; it < 4; ++it) {
      cudaEventRecord(e0, s0);
      c.run(s0);
      if (cudaGetLastError() != cudaSuccess) { best = -1; break; }  // e.g. 1024 threads x too many registers
      cudaEventRecord(e1, s0);
      cudaError_t e = cudaStreamSynchronize(s0);
      if (e != cudaSuccess) { printf("%s: %s\n", c.name.c_str(), cudaGetErrorString(e)); return 1; }
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (it) best = std::min(best, ms);
    }
    if (best < 0) printf("%-52s launch failed (skipped)\n", c.name.c_str());
    else printf("%-52s %7.1f GB/s\n", c.name.c_str(), bytes / (best * 1e-3) / 1e9);
  }
  // the same copy issued from the DESTINATION's stream (pull, if the driver uses that device's engine)
  {
    cudaSetDevice(1);
    cudaEvent_t f0, f1; cudaEventCreate(&f0); cudaEventCreate(&f1);
    float best = 1e30f;
    for (int it = 0; it < 4; ++it) {
      cudaEventRecord(f0, s1);
      cudaMemcpyPeerAsync(d1, 1, d0, 0, bytes, s1);
      cudaEventRecord(f1, s1);
      cudaStreamSynchronize(s1);
      float ms; cudaEventElapsedTime(&ms, f0, f1);
      if (it) best = std::min(best, ms);
    }
    printf("%-52s %7.1f GB/s\n", "copy engine, dst stream", bytes / (best * 1e-3) / 1e9);
  }
  return 0;
}
