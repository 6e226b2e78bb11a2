// nvml_lite.cu — a handful of NVML queries through dlopen("libnvidia-ml.so.1"), for two things
// the probe needs that CUDA cannot tell it:
//   * is anybody else computing on this GPU right now (a 16 GiB, full-SM probe
//     perturbs tenants — the plugin gates the periodic re-probe on this), and
//   * clocks / power / throttle reasons beside a measurement.
// The reference reaches NVML through go-nvml (go.mod:8), which dlopens the same library
// (device/device.go:37-181 are its call sites); this file does not replace any of that.
// No link-time dependency: without the library every query reports available = 0.
#include <dlfcn.h>
#include <unistd.h>

#include <cstring>
#include <mutex>

#include "engine.h"

namespace {

typedef int nvmlReturn_t;
typedef void* nvmlDevice_t;
struct nvmlMemory_t { unsigned long long total, free, used; };
struct nvmlProcessInfo_t { unsigned int pid; unsigned long long usedGpuMemory; unsigned int gpuInstanceId, computeInstanceId; };

struct Nvml {
  void* h = nullptr;
  bool ok = false;
  nvmlReturn_t (*Init)(void) = nullptr;
  nvmlReturn_t (*GetHandleByUUID)(const char*, nvmlDevice_t*) = nullptr;
  nvmlReturn_t (*GetClockInfo)(nvmlDevice_t, int, unsigned int*) = nullptr;
  nvmlReturn_t (*GetMaxClockInfo)(nvmlDevice_t, int, unsigned int*) = nullptr;
  nvmlReturn_t (*GetPowerUsage)(nvmlDevice_t, unsigned int*) = nullptr;
  nvmlReturn_t (*GetTemperature)(nvmlDevice_t, int, unsigned int*) = nullptr;
  nvmlReturn_t (*GetThrottle)(nvmlDevice_t, unsigned long long*) = nullptr;
  nvmlReturn_t (*GetProcs)(nvmlDevice_t, unsigned int*, nvmlProcessInfo_t*) = nullptr;
  nvmlReturn_t (*GetMemoryInfo)(nvmlDevice_t, nvmlMemory_t*) = nullptr;
};

Nvml* nvml() {
  static Nvml n;
  static std::once_flag once;
  std::call_once(once, [] {
    n.h = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!n.h) return;
    auto sym = [&](const char* s) { return dlsym(n.h, s); };
    n.Init = reinterpret_cast<decltype(n.Init)>(sym("nvmlInit_v2"));
    n.GetHandleByUUID = reinterpret_cast<decltype(n.GetHandleByUUID)>(sym("nvmlDeviceGetHandleByUUID"));
    n.GetClockInfo = reinterpret_cast<decltype(n.GetClockInfo)>(sym("nvmlDeviceGetClockInfo"));
    n.GetMaxClockInfo = reinterpret_cast<decltype(n.GetMaxClockInfo)>(sym("nvmlDeviceGetMaxClockInfo"));
    n.GetPowerUsage = reinterpret_cast<decltype(n.GetPowerUsage)>(sym("nvmlDeviceGetPowerUsage"));
    n.GetTemperature = reinterpret_cast<decltype(n.GetTemperature)>(sym("nvmlDeviceGetTemperature"));
    n.GetThrottle = reinterpret_cast<decltype(n.GetThrottle)>(sym("nvmlDeviceGetCurrentClocksThrottleReasons"));
    n.GetProcs = reinterpret_cast<decltype(n.GetProcs)>(sym("nvmlDeviceGetComputeRunningProcesses_v3"));
    if (!n.GetProcs) n.GetProcs = reinterpret_cast<decltype(n.GetProcs)>(sym("nvmlDeviceGetComputeRunningProcesses_v2"));
    n.GetMemoryInfo = reinterpret_cast<decltype(n.GetMemoryInfo)>(sym("nvmlDeviceGetMemoryInfo"));
    n.ok = n.Init && n.GetHandleByUUID && n.Init() == 0;
  });
  return &n;
}

// does THIS process hold an active primary context on the device (ours, or anybody else's in the process)?
bool self_has_context(int ordinal) {
  typedef int (*GetDevFn)(int*, int);
  typedef int (*StateFn)(int, unsigned int*, int*);
  static GetDevFn get_dev = nullptr;
  static StateFn state = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuDeviceGet", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      get_dev = reinterpret_cast<GetDevFn>(p);
    if (cudaGetDriverEntryPoint("cuDevicePrimaryCtxGetState", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      state = reinterpret_cast<StateFn>(p);
    (void)cudaGetLastError();
  });
  if (!get_dev || !state) return false;
  int dev = 0, active = 0;
  unsigned int flags = 0;
  if (get_dev(&dev, ordinal) != 0 || state(dev, &flags, &active) != 0) return false;
  return active != 0;
}

}  // namespace

extern "C" int gp_nvml_snapshot(gp_ctx* ctx, const char* uuid, gp_nvml_info* out) {
  using namespace gp;
  if (!ctx || !out) { set_error("null argument"); return GP_ERR_INVALID; }
  Device* d = find_device(ctx, uuid);
  if (!d) { set_error("device '%s' is not visible to CUDA", uuid ? uuid : "(null)"); return GP_ERR_NO_DEVICE; }
  memset(out, 0, sizeof(*out));
  out->struct_size = (uint32_t)sizeof(*out);
  Nvml* n = nvml();
  if (!n->ok) return GP_OK;  // available = 0: not an error, the caller decides what that means
  nvmlDevice_t h = nullptr;
  if (n->GetHandleByUUID(d->uuid, &h) != 0 || !h) return GP_OK;
  out->available = 1;
  unsigned int v = 0;
  if (n->GetClockInfo && n->GetClockInfo(h, 1, &v) == 0) out->sm_clock_mhz = v;
  if (n->GetMaxClockInfo && n->GetMaxClockInfo(h, 1, &v) == 0) out->sm_clock_max_mhz = v;
  if (n->GetClockInfo && n->GetClockInfo(h, 2, &v) == 0) out->mem_clock_mhz = v;
  if (n->GetPowerUsage && n->GetPowerUsage(h, &v) == 0) out->power_mw = v;
  if (n->GetTemperature && n->GetTemperature(h, 0, &v) == 0) out->temperature_c = v;
  unsigned long long r = 0;
  if (n->GetThrottle && n->GetThrottle(h, &r) == 0) out->throttle_reasons = r;
  nvmlMemory_t m{};
  if (n->GetMemoryInfo && n->GetMemoryInfo(h, &m) == 0) { out->mem_used_bytes = m.used; out->mem_total_bytes = m.total; }
  if (n->GetProcs) {
    nvmlProcessInfo_t infos[64];
    unsigned int count = 64;
    const nvmlReturn_t rc = n->GetProcs(h, &count, infos);
    if (rc == 0) {
      // "Other" must not count this process.  NVML reports pids of the HOST pid namespace; in a pod without
      // hostPID getpid() is a different number and the caller would see itself as a tenant on every GPU it has
      // ever touched (and never grade it again).  So: drop our pid if it is in the list; if it is not, but this
      // process holds an active primary context on the device, exactly one of the listed processes is us.
      const unsigned int self = (unsigned int)getpid();
      bool found_self = false;
      for (unsigned int i = 0; i < count && i < 64; ++i) found_self = found_self || infos[i].pid == self;
      unsigned int others = count;
      if (found_self || (count > 0 && self_has_context(d->ordinal))) others -= 1;
      out->other_compute_procs = others;
    } else if (rc == 7 /* NVML_ERROR_INSUFFICIENT_SIZE */) {
      out->other_compute_procs = count;
    }
  }
  return GP_OK;
}
