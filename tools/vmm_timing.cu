// vmm_timing — what the arena costs: cudaMalloc/cudaFree against cuMemCreate/Map/SetAccess/Unmap/Release
// at several chunk sizes, alone and with every visible GPU doing the same at once (one thread per GPU,
// the shape of gp_probe_all).  Development tool; results in profiles/arena_timing_r02.txt.
//   nvcc -O2 -std=c++17 -o tools/vmm_timing tools/vmm_timing.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <thread>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define CK(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { printf("%s -> %d\n", #x, (int)r_); return; } } while (0)

static void vmm_cycle(int dev, size_t total, size_t chunk, bool one_unmap, double* t_alloc, double* t_free, double* t_parts) {
  cudaSetDevice(dev);
  cudaFree(0);
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  CUdeviceptr base = 0;
  CK(cuMemAddressReserve(&base, total, 1ull << 30, 0, 0));
  std::vector<CUmemGenericAllocationHandle> hs;
  double t0 = now(), tc = 0, tm = 0, ta = 0;
  for (size_t off = 0; off < total; off += chunk) {
    CUmemGenericAllocationHandle h;
    double a = now();
    CK(cuMemCreate(&h, chunk, &prop, 0));
    double b = now();
    CK(cuMemMap(base + off, chunk, 0, h, 0));
    double c = now();
    CUmemAccessDesc acc{};
    acc.location = prop.location;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CK(cuMemSetAccess(base + off, chunk, &acc, 1));
    double d = now();
    tc += b - a; tm += c - b; ta += d - c;
    hs.push_back(h);
  }
  *t_alloc = now() - t0;
  cudaMemset((void*)base, 1, 1 << 20);
  cudaDeviceSynchronize();
  double t1 = now(), tu = 0, tr = 0;
  if (one_unmap) { double a = now(); CK(cuMemUnmap(base, total)); tu += now() - a; }
  size_t off = 0;
  for (auto h : hs) {
    double a = now();
    if (!one_unmap) CK(cuMemUnmap(base + off, chunk));
    double b = now();
    CK(cuMemRelease(h));
    double c = now();
    tu += b - a; tr += c - b;
    off += chunk;
  }
  *t_free = now() - t1;
  CK(cuMemAddressFree(base, total));
  t_parts[0] = tc; t_parts[1] = tm; t_parts[2] = ta; t_parts[3] = tu; t_parts[4] = tr;
}

static void malloc_cycle(int dev, size_t total, double* t_alloc, double* t_free) {
  cudaSetDevice(dev);
  cudaFree(0);
  void* p = nullptr;
  double t0 = now();
  cudaMalloc(&p, total);
  *t_alloc = now() - t0;
  cudaMemset(p, 1, 1 << 20);
  cudaDeviceSynchronize();
  double t1 = now();
  cudaFree(p);
  *t_free = now() - t1;
}

int main() {
  int n = 0;
  cudaGetDeviceCount(&n);
  cuInit(0);
  const size_t total = 48ull << 30;
  for (int ng : {1, n}) {
    if (ng == n && n == 1) break;
    printf("== %d GPU(s) at once, 48 GiB each\n", ng);
    for (int rep = 0; rep < 2; ++rep) {
      std::vector<double> a(ng), f(ng);
      std::vector<std::thread> th;
      for (int d = 0; d < ng; ++d) th.emplace_back([&, d] { malloc_cycle(d, total, &a[d], &f[d]); });
      for (auto& t : th) t.join();
      double ma = 0, mf = 0;
      for (int d = 0; d < ng; ++d) { ma = std::max(ma, a[d]); mf = std::max(mf, f[d]); }
      printf("cudaMalloc/cudaFree          rep %d: alloc %7.1f ms  free %7.1f ms (max over GPUs)\n", rep, ma * 1e3, mf * 1e3);
    }
    for (size_t chunk : {256ull << 20, 1ull << 30, 4ull << 30, 16ull << 30, 48ull << 30})
      for (int one : {0, 1}) {
        std::vector<double> a(ng), f(ng);
        std::vector<std::vector<double>> parts(ng, std::vector<double>(5));
        std::vector<std::thread> th;
        for (int d = 0; d < ng; ++d) th.emplace_back([&, d] { vmm_cycle(d, total, chunk, one != 0, &a[d], &f[d], parts[d].data()); });
        for (auto& t : th) t.join();
        double ma = 0, mf = 0;
        for (int d = 0; d < ng; ++d) { ma = std::max(ma, a[d]); mf = std::max(mf, f[d]); }
        printf("VMM chunk %6zu MiB %s: alloc %7.1f ms  free %7.1f ms | gpu0 create %.1f map %.1f access %.1f unmap %.1f release %.1f ms\n",
               chunk >> 20, one ? "one unmap " : "per-chunk ", ma * 1e3, mf * 1e3, parts[0][0] * 1e3, parts[0][1] * 1e3, parts[0][2] * 1e3,
               parts[0][3] * 1e3, parts[0][4] * 1e3);
      }
  }
  return 0;
}
