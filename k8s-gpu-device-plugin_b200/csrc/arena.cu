// arena.cu — the per-device arena every probe carves its arrays from.
//
// No reference counterpart (the reference never allocates device memory).  Two backends:
//
//   VMM (AUTO's choice with keep_arena = 1): one virtual range per device is reserved on first use
//   (cuMemAddressReserve; virtual space is free) and physical chunks (cuMemCreate, 4 GiB each plus one
//   tail chunk) are mapped from its base upwards as the arena grows.  Growing never moves what is
//   already mapped, so a kept arena (the periodic re-probe of an idle device) costs nothing inside the
//   probe call and a larger request only maps the difference.  gp_trim / gp_trim_device unmap and
//   release the physical memory (pods get it back) but keep the reservation.
//
//   cudaMalloc (AUTO's choice with keep_arena = 0): one allocation per size change.  Measured
//   (profiles/arena_timing_r02.txt): it is the cheapest way to get and return 48 GiB cold — 5 ms +
//   11..19 ms against 12 + 35 ms for VMM in 4 GiB chunks; the VMM calls cost 0.2 - 1 ms EACH.
//
// Either way the driver serialises memory management across threads and processes (eight processes
// freeing 48 GiB at once: 225 ms, SCALE_r01), so what cannot be avoided is taken off the caller's
// path: a cold probe hands its memory to a reaper thread and returns (free_arena_deferred /
// free_buffers_deferred; the next call on the device joins it first), and a cold gp_probe_device does
// not use this arena at all: it takes three cudaMalloc slabs, allocated in order on a provider thread
// behind the probes that do not need them yet (engine.cu, probe_device_impl).  Device::alloc_s /
// free_s report the host time a call really spent blocked on the allocator.
#include <cuda.h>

#include <chrono>

#include "engine.h"

namespace gp {

namespace {

struct VmmApi {
  bool ok = false;
  CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*GetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
};

template <class F>
bool entry(const char* name, F* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    (void)cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(p);
  return true;
}

const VmmApi* vmm_api() {
  static VmmApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    api.ok = entry("cuMemAddressReserve", &api.AddressReserve) && entry("cuMemAddressFree", &api.AddressFree) &&
             entry("cuMemCreate", &api.Create) && entry("cuMemRelease", &api.Release) &&
             entry("cuMemMap", &api.Map) && entry("cuMemUnmap", &api.Unmap) &&
             entry("cuMemSetAccess", &api.SetAccess) &&
             entry("cuMemGetAllocationGranularity", &api.GetGranularity);
  });
  return &api;
}

constexpr size_t kChunkBytes = 4ull << 30;
constexpr size_t kVaAlign = 1ull << 30;

int cu_fail(CUresult r, const char* what) {
  set_error("%s failed: CUresult %d", what, (int)r);
  return r == CUDA_ERROR_OUT_OF_MEMORY ? GP_ERR_NOMEM : GP_ERR_CUDA;
}

bool device_supports_vmm(int ordinal) {
  CUresult (*GetDev)(CUdevice*, int) = nullptr;
  CUresult (*GetAttr)(int*, CUdevice_attribute, CUdevice) = nullptr;
  if (!entry("cuDeviceGet", &GetDev) || !entry("cuDeviceGetAttribute", &GetAttr)) return false;
  CUdevice dev;
  int vmm = 0;
  if (GetDev(&dev, ordinal) != CUDA_SUCCESS) return false;
  if (GetAttr(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev) != CUDA_SUCCESS) return false;
  return vmm != 0;
}

CUmemAccessDesc access_for(int ordinal) {
  CUmemAccessDesc a{};
  a.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  a.location.id = ordinal;
  a.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  return a;
}

// unmap + release everything mapped above `keep_bytes` (a chunk boundary); VMM backend only.  One
// cuMemUnmap covers all the chunks being dropped (measured: 3 ms against 17 ms chunk by chunk).
void vmm_shrink(Device* d, size_t keep_bytes) {
  const VmmApi* api = vmm_api();
  size_t off = d->arena_bytes;
  size_t first = d->chunks.size();
  while (first > 0 && off > keep_bytes) { off -= d->chunks[first - 1].bytes; --first; }
  if (first < d->chunks.size()) {
    api->Unmap((CUdeviceptr)(d->va_base + off), d->arena_bytes - off);
    for (size_t i = first; i < d->chunks.size(); ++i) api->Release((CUmemGenericAllocationHandle)d->chunks[i].handle);
    d->chunks.resize(first);
  }
  d->arena_bytes = off;
  if (off == 0) { d->arena = nullptr; d->peer_mask = 0; }
}

int vmm_grow(Ctx* ctx, Device* d, size_t bytes) {
  (void)ctx;
  const VmmApi* api = vmm_api();
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = d->ordinal;
  size_t gran = 0;
  CUresult r = api->GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS || gran == 0) return cu_fail(r, "cuMemGetAllocationGranularity");
  if (!d->va_base) {
    // everything the device could ever hold, rounded up: virtual space costs nothing
    size_t va = ((size_t)d->total_mem + 2 * kChunkBytes - 1) / kChunkBytes * kChunkBytes;
    CUdeviceptr base = 0;
    r = api->AddressReserve(&base, va, kVaAlign, 0, 0);
    if (r != CUDA_SUCCESS) return cu_fail(r, "cuMemAddressReserve");
    d->va_base = (unsigned long long)base;
    d->va_bytes = va;
  }
  if (bytes > d->va_bytes) { set_error("arena request %zu exceeds the device's address range %zu", bytes, d->va_bytes); return GP_ERR_NOMEM; }
  const size_t old_bytes = d->arena_bytes;
  size_t off = old_bytes;
  while (off < bytes) {
    size_t want = std::min(kChunkBytes, bytes - off);
    want = (want + gran - 1) / gran * gran;
    CUmemGenericAllocationHandle h = 0;
    r = api->Create(&h, want, &prop, 0);
    if (r != CUDA_SUCCESS) { d->arena_bytes = off; vmm_shrink(d, old_bytes); return cu_fail(r, "cuMemCreate"); }
    r = api->Map((CUdeviceptr)(d->va_base + off), want, 0, h, 0);
    if (r != CUDA_SUCCESS) { api->Release(h); d->arena_bytes = off; vmm_shrink(d, old_bytes); return cu_fail(r, "cuMemMap"); }
    d->chunks.push_back({(unsigned long long)h, want});
    off += want;
  }
  // access for the owner and for every peer that already sees the lower part
  std::vector<CUmemAccessDesc> acc;
  acc.push_back(access_for(d->ordinal));
  for (int i = 0; i < 64; ++i)
    if (i != d->ordinal && (d->peer_mask >> i & 1)) acc.push_back(access_for(i));
  r = api->SetAccess((CUdeviceptr)(d->va_base + old_bytes), off - old_bytes, acc.data(), acc.size());
  if (r != CUDA_SUCCESS) { d->arena_bytes = off; vmm_shrink(d, old_bytes); return cu_fail(r, "cuMemSetAccess"); }
  d->arena = reinterpret_cast<uint8_t*>(d->va_base);
  d->arena_bytes = off;
  d->arena_is_vmm = true;
  return GP_OK;
}

bool want_vmm(const Ctx* ctx, const Device* d) {
  if (ctx->cfg.arena_backend == GP_ARENA_MALLOC) return false;
  if (ctx->cfg.arena_backend == GP_ARENA_AUTO && !ctx->cfg.keep_arena) return false;  // cold every call: cudaMalloc is cheaper
  static bool supported[64] = {};
  static bool known[64] = {};
  if (!vmm_api()->ok || d->ordinal < 0 || d->ordinal >= 64) return false;
  if (!known[d->ordinal]) { supported[d->ordinal] = device_supports_vmm(d->ordinal); known[d->ordinal] = true; }
  return supported[d->ordinal];
}

struct Stopwatch {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  double seconds() const { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
};

}  // namespace

uint64_t cfg_arena_cap(const Ctx* c) {
  const uint64_t hbm = c->cfg.hbm_bytes ? c->cfg.hbm_bytes : (16ull << 30);
  return c->cfg.arena_cap_bytes ? c->cfg.arena_cap_bytes : 5 * hbm + (1ull << 30);
}

void arena_join_reaper(Device* d) {
  if (d->reaper.joinable()) d->reaper.join();
}

// Release the arena on a background thread and return at once: the caller's results are complete, what is left
// is handing memory back, and the driver takes tens to hundreds of milliseconds over that when several GPUs do it
// together.  The memory is back with the driver when the thread ends; anything on this device that needs the
// allocator again (ensure_arena, free_arena, trim, reset, shutdown) joins it first.
void free_arena_deferred(Device* d) {
  arena_join_reaper(d);
  if (!d->arena) { d->arena_bytes = 0; d->triad_operands_bytes = 0; return; }
  uint8_t* arena = d->arena;
  const bool vmm = d->arena_is_vmm;
  const unsigned long long va = d->va_base;
  const size_t bytes = d->arena_bytes;
  std::vector<Device::Chunk> chunks;
  chunks.swap(d->chunks);
  const int ordinal = d->ordinal;
  d->arena = nullptr;
  d->arena_bytes = 0;
  d->peer_mask = 0;
  d->triad_operands_bytes = 0;
  d->reaper = std::thread([=] {
    cudaSetDevice(ordinal);
    if (vmm) {
      const VmmApi* api = vmm_api();
      api->Unmap((CUdeviceptr)va, bytes);
      for (const auto& c : chunks) api->Release((CUmemGenericAllocationHandle)c.handle);
    } else {
      cudaFree(arena);
    }
    (void)cudaGetLastError();
  });
}

// the same for plain cudaMalloc buffers that are not the arena (the cold probe's slabs)
void free_buffers_deferred(Device* d, const std::vector<void*>& ptrs) {
  arena_join_reaper(d);
  if (ptrs.empty()) return;
  const int ordinal = d->ordinal;
  d->reaper = std::thread([ptrs, ordinal] {
    cudaSetDevice(ordinal);
    for (void* p : ptrs) cudaFree(p);
    (void)cudaGetLastError();
  });
}

void free_arena(Device* d) {
  arena_join_reaper(d);
  if (!d->arena) { d->arena_bytes = 0; d->triad_operands_bytes = 0; return; }
  Stopwatch sw;
  cudaSetDevice(d->ordinal);
  if (d->arena_is_vmm) {
    vmm_shrink(d, 0);
  } else {
    cudaFree(d->arena);
  }
  d->arena = nullptr;
  d->arena_bytes = 0;
  d->peer_mask = 0;
  d->triad_operands_bytes = 0;
  d->free_s += sw.seconds();
}

void release_arena_va(Device* d) {
  free_arena(d);
  if (d->va_base && vmm_api()->ok) vmm_api()->AddressFree((CUdeviceptr)d->va_base, d->va_bytes);
  d->va_base = 0;
  d->va_bytes = 0;
  d->arena_is_vmm = false;
}

int ensure_arena(Ctx* ctx, Device* d, size_t bytes, bool ignore_cap) {
  if (bytes <= d->arena_bytes) return GP_OK;
  if (!ignore_cap && bytes > cfg_arena_cap(ctx)) {
    set_error("arena request %zu exceeds cap %llu", bytes, (unsigned long long)cfg_arena_cap(ctx));
    return GP_ERR_NOMEM;
  }
  Stopwatch sw;
  arena_join_reaper(d);  // memory on its way back to the driver must have arrived before asking for more
  const bool vmm = want_vmm(ctx, d);
  if (d->arena && d->arena_is_vmm != vmm) free_arena(d);
  if (!vmm) free_arena(d);  // cudaMalloc cannot grow in place
  size_t free_b = 0, total_b = 0;
  cudaError_t e = cudaMemGetInfo(&free_b, &total_b);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemGetInfo", nullptr);
  int rc = GP_OK;
  if (vmm) {
    if (bytes - d->arena_bytes > free_b) {
      set_error("arena request %zu exceeds free device memory %zu (+%zu already mapped)", bytes, free_b, d->arena_bytes);
      rc = GP_ERR_NOMEM;
    } else {
      rc = vmm_grow(ctx, d, bytes);
    }
  } else {
    if (bytes > free_b) {
      set_error("arena request %zu exceeds free device memory %zu", bytes, free_b);
      rc = GP_ERR_NOMEM;
    } else {
      void* p = nullptr;
      e = cudaMalloc(&p, bytes);
      if (e != cudaSuccess) rc = cuda_fail(e, "cudaMalloc(arena)", nullptr);
      else {
        d->arena = reinterpret_cast<uint8_t*>(p);
        d->arena_bytes = bytes;
        d->arena_is_vmm = false;
      }
    }
  }
  d->alloc_s += sw.seconds();
  return rc;
}

int arena_grant_peers(Device* d, unsigned long long ordinal_mask) {
  if (!d->arena || !d->arena_is_vmm) return GP_OK;
  const unsigned long long add = ordinal_mask & ~d->peer_mask & ~(1ull << d->ordinal);
  if (!add) return GP_OK;
  std::vector<CUmemAccessDesc> acc;
  for (int i = 0; i < 64; ++i)
    if (add >> i & 1) acc.push_back(access_for(i));
  const CUresult r = vmm_api()->SetAccess((CUdeviceptr)d->va_base, d->arena_bytes, acc.data(), acc.size());
  if (r != CUDA_SUCCESS) return cu_fail(r, "cuMemSetAccess(peers)");
  d->peer_mask |= add;
  return GP_OK;
}

}  // namespace gp
