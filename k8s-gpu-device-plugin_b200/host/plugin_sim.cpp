// plugin_sim — walks the reference's start-up path (main.go:74-131 -> plugin/manager.go:56-99)
// with the probe wired in, on real GPUs, and prints one JSON document:
//   NewDeviceMap -> GradeDevices -> ListAndWatch first list -> GET /probe -> GET /metrics
// There is no kubelet, gRPC or HTTP server here (out of scope); the handlers are called
// directly.  Used by tests/test_host_gpu.py.
//   plugin_sim [--hbm-mib N] [--pattern-mib N] [--wo N] [--gemm N] [--mask M] [--slice BIN] [--min-copy G]
//              [--scenario manager] [--fabric-mib N] [--min-p2p G]
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <thread>

#include "manager.hpp"
#include "router.hpp"

using namespace plugin;

// deviceInfo backed by the probe library's enumeration (the reference's nvmlDevice,
// device/device.go:21-181, asks NVML; CUDA reports the same UUIDs)
struct CudaDeviceInfo : deviceInfo {
  gp_device_result id;
  std::pair<std::string, Error> GetUUID() override { return {id.uuid, std::nullopt}; }
  std::pair<std::vector<std::string>, Error> GetPaths() override {
    return {{"/dev/nvidia" + std::to_string(id.ordinal)}, std::nullopt};
  }
  std::tuple<bool, int, Error> GetNumaNode() override { return {false, 0, std::nullopt}; }
  std::pair<uint64_t, Error> GetTotalMemory() override { return {id.total_mem, std::nullopt}; }
  std::pair<std::string, Error> GetComputeCapability() override {
    return {std::to_string(id.cc_major) + "." + std::to_string(id.cc_minor), std::nullopt};
  }
};

static std::string listJson(const std::vector<PluginDevice>& l) {
  std::string s = "[";
  for (size_t i = 0; i < l.size(); ++i)
    s += std::string(i ? "," : "") + "{\"ID\":\"" + l[i].ID + "\",\"health\":\"" + l[i].Health + "\"}";
  return s + "]";
}

// --scenario manager: the PluginManager path on real GPUs — loadPlugins grades every GPU and then the fabric
// (P2P matrix + all-reduce) in one go, the matrix feeds the gauges and GetPreferredAllocation; then the
// isolation path: a child that faults, a child that never answers (killed at the deadline), and the way back.
// --assume-idle 1: the harness that starts this binary (pytest, with its own CUDA context on GPU 0) is itself a
// "tenant" in NVML's eyes; a test that wants every GPU graded tells the backend to treat them as idle
struct AssumeIdleBackend : GpuProbeBackend {
  using GpuProbeBackend::GpuProbeBackend;
  bool assume_idle = false;
  int OtherComputeProcs(const std::string& u) override { return assume_idle ? 0 : GpuProbeBackend::OtherComputeProcs(u); }
};

static int manager_scenario(gp_config cfg, uint32_t mask, const std::string& slice_bin, bool assume_idle) {
  cfg.keep_arena = 1;
  AssumeIdleBackend be(&cfg, slice_bin);
  be.assume_idle = assume_idle;
  if (be.init_rc()) { printf("{\"error\":\"gp_init: %s\"}\n", be.init_error().c_str()); return 1; }
  const int n = gp_device_count(be.ctx());
  std::vector<std::shared_ptr<CudaDeviceInfo>> infos;
  for (int i = 0; i < n; ++i) {
    auto info = std::make_shared<CudaDeviceInfo>();
    char uuid[64];
    gp_device_uuid(be.ctx(), i, uuid, sizeof(uuid));
    gp_probe_device(be.ctx(), uuid, 0, &info->id);
    infos.push_back(info);
  }
  auto enumerate = [&] {
    std::vector<GpuNode> nodes;
    for (int i = 0; i < n; ++i) nodes.push_back({i, "NVIDIA B200", false, infos[(size_t)i], {}});
    return nodes;
  };
  ProbeMetrics metrics;
  PluginManager mgr(MigStrategyNone, {Resource::New("*", "gpu")}, enumerate, &be, &metrics, mask);
  if (auto e = mgr.loadPlugins()) { printf("{\"error\":\"%s\"}\n", jsonEscape(*e).c_str()); return 1; }
  const gp_fabric_result& f = mgr.fabric();
  std::vector<std::string> all_ids = mgr.devices().m["nvidia.com/gpu"].GetIDs();
  std::string pref = "[";
  if (n >= 2) {
    auto [pick, pe] = mgr.PreferredAllocation("nvidia.com/gpu", all_ids, {}, std::min(n, 2));
    for (size_t i = 0; i < pick.size(); ++i) pref += std::string(i ? "," : "") + "\"" + pick[i] + "\"";
  }
  pref += "]";
  auto plug = mgr.plugins()[0];
  std::vector<std::string> lists;
  std::thread law([&] { plug->ListAndWatch([&](const std::vector<PluginDevice>& l) { lists.push_back(listJson(l)); return true; }); });
  auto health0 = [&] { return std::string(VerdictHealthy(*mgr.devices().m["nvidia.com/gpu"].GetByID(all_ids[0])) ? "Healthy" : "Unhealthy"); };
  std::string steps = "[";
  auto step = [&](const char* name) {
    auto [graded, err] = mgr.Reprobe(all_ids[0]);
    const Device& d = *mgr.devices().m["nvidia.com/gpu"].GetByID(all_ids[0]);
    steps += std::string(steps.size() > 1 ? "," : "") + "{\"step\":\"" + name + "\",\"verdict\":\"" + health0() + "\",\"error\":" +
             std::to_string(d.Probe ? d.Probe->error : 0) + ",\"graded\":" + std::to_string(graded.size()) + "}";
  };
  int rc = 0;
  if (!slice_bin.empty()) {
    mgr.SetIsolation(true);
    step("isolated");                                      // fresh child, healthy
    be.SetChildExtraArgs({"--test-cuda-error", "700"});    // the child's device reports a sticky fault
    step("child_fault");
    be.SetChildExtraArgs({});
    step("after_fault");                                   // next child: healthy again, parent untouched
    be.SetChildExtraArgs({"--test-hang"});                 // a child that never answers
    be.SetChildTimeout(std::chrono::milliseconds(1500));
    step("child_hang");
    be.SetChildExtraArgs({});
    be.SetChildTimeout(std::chrono::milliseconds(120000));
    step("after_hang");
    mgr.SetIsolation(false);
    step("in_process");                                    // and the plugin's own context still works
  }
  plug->Stop();
  law.join();
  steps += "]";
  printf("{\"devices\":%d,\"unhealthy\":%d,\"fabric\":{\"ngpus\":%u,\"healthy\":%d,\"data_ok\":%u,\"failed\":%u,\"spread\":%.6f,"
         "\"read_min\":%.1f,\"read_max\":%.1f,\"write_min\":%.1f,\"write_max\":%.1f,\"worst\":[%d,%d],\"allreduce_busbw\":%.1f,"
         "\"allreduce_error\":%d,\"seconds\":%.3f},\"preferred\":%s,\"steps\":%s,\"isolated_probes\":%d,\"lists\":[",
         n, mgr.unhealthyAtLoad(), f.ngpus, f.healthy, f.data_ok, f.failed, f.spread, f.read_min, f.read_max, f.write_min, f.write_max,
         f.worst_src, f.worst_dst, f.allreduce_busbw_gbps, f.allreduce_error, f.duration_s, pref.c_str(), steps.c_str(), mgr.isolatedProbes());
  for (size_t i = 0; i < lists.size(); ++i) printf("%s%s", i ? "," : "", lists[i].c_str());
  printf("],\"log\":[");
  auto log = mgr.log();
  for (size_t i = 0; i < log.size(); ++i) printf("%s\"%s\"", i ? "," : "", jsonEscape(log[i]).c_str());
  printf("],\"busy_skipped\":%d,\"metrics\":\"%s\"}\n", mgr.skippedBusy(), jsonEscape(metrics.Render()).c_str());
  fflush(stdout);
  return rc;
}

// a crash in the harness must say where: print the stack and leave (no core files on the GPU boxes)
static void on_fatal_signal(int sig) {
  void* frames[48];
  const int n = backtrace(frames, 48);
  dprintf(STDERR_FILENO, "plugin_sim: fatal signal %d, stack:\n", sig);
  backtrace_symbols_fd(frames, n, STDERR_FILENO);
  _exit(128 + sig);
}

int main(int argc, char** argv) {
  signal(SIGSEGV, on_fatal_signal);
  signal(SIGABRT, on_fatal_signal);
  gp_config cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.struct_size = sizeof(cfg);
  cfg.abi_version = GP_ABI_VERSION;
  cfg.hbm_bytes = 256ull << 20;
  cfg.pattern_bytes = 64ull << 20;
  cfg.wo_passes = 4;
  cfg.gemm_mnk = 1024;
  cfg.iters = 3;
  cfg.warmup = 1;
  uint32_t mask = GP_PROBE_ALL_SINGLE;
  std::string slice_bin, scenario;
  bool assume_idle = false;
  for (int i = 1; i + 1 < argc; i += 2) {
    const std::string a = argv[i];
    const char* v = argv[i + 1];
    if (a == "--hbm-mib") cfg.hbm_bytes = strtoull(v, nullptr, 10) << 20;
    else if (a == "--pattern-mib") cfg.pattern_bytes = strtoull(v, nullptr, 10) << 20;
    else if (a == "--wo") cfg.wo_passes = (uint32_t)atoi(v);
    else if (a == "--gemm") cfg.gemm_mnk = (uint32_t)atoi(v);
    else if (a == "--mask") mask = (uint32_t)atoi(v);
    else if (a == "--slice") slice_bin = v;
    else if (a == "--min-copy") cfg.min_copy_gbps = atof(v);
    else if (a == "--scenario") scenario = v;
    else if (a == "--fabric-mib") cfg.fabric_bytes = strtoull(v, nullptr, 10) << 20;
    else if (a == "--min-p2p") cfg.min_p2p_gbps = atof(v);
    else if (a == "--assume-idle") assume_idle = atoi(v) != 0;
  }
  if (scenario == "manager") return manager_scenario(cfg, mask, slice_bin, assume_idle);
  GpuProbeBackend be(&cfg, slice_bin);
  if (be.init_rc()) {
    printf("{\"error\":\"gp_init: %s\"}\n", be.init_error().c_str());
    return 1;
  }
  // NewDeviceMap (device_map.go:24-75): one entry per visible GPU under nvidia.com/gpu
  DeviceMap dm;
  const int n = gp_device_count(be.ctx());
  for (int i = 0; i < n; ++i) {
    CudaDeviceInfo info;
    char uuid[64];
    gp_device_uuid(be.ctx(), i, uuid, sizeof(uuid));
    gp_probe_device(be.ctx(), uuid, 0, &info.id);  // mask 0: identity only
    if (!matchResourcePattern("*", "NVIDIA B200")) return 2;
    if (auto e = dm.setEntry("nvidia.com/gpu", std::to_string(i), info)) { printf("{\"error\":\"%s\"}\n", e->c_str()); return 1; }
    if (!slice_bin.empty() && i == 0) {
      // no MIG on this box: exercise the child-process path by naming GPU 0 as slice "0:0" of a
      // second resource; CUDA_VISIBLE_DEVICES accepts a GPU uuid exactly as it accepts a MIG uuid
      struct SliceInfo : CudaDeviceInfo {
        std::pair<std::string, Error> GetUUID() override { return {std::string(id.uuid) + "::1", std::nullopt}; }
      } si;
      si.id = info.id;
      if (auto e = dm.setEntry("nvidia.com/mig-sim", "0:0", si)) { printf("{\"error\":\"%s\"}\n", e->c_str()); return 1; }
    }
  }
  std::vector<std::string> log;
  const int unhealthy = GradeDevices(dm, be, mask, &log);
  ProbeMetrics metrics;
  for (auto& d : dm.All()) metrics.Update(*d);

  NvidiaDevicePlugin plug("nvidia.com/gpu", dm.m["nvidia.com/gpu"]);
  std::vector<std::string> lists;
  std::thread law([&] {
    plug.ListAndWatch([&](const std::vector<PluginDevice>& l) { lists.push_back(listJson(l)); return true; });
  });
  API api(&dm, &be, &metrics, mask, [&](const std::string& id, bool h) { plug.Report(id, h); }, nullptr);
  const Response probe = api.Probe();
  // a fault report and its recovery must each produce exactly one resend
  const std::string first_id = dm.m["nvidia.com/gpu"].GetIDs().front();
  plug.Report(first_id, false);
  plug.Report(first_id, false);
  plug.Report(first_id, true);
  plug.Stop();
  law.join();
  const Response unknown = api.Probe("GPU-00000000-0000-0000-0000-000000000000");
  const std::string m = api.Metrics().body;

  printf("{\"devices\":%d,\"unhealthy\":%d,\"lists\":[", n, unhealthy);
  for (size_t i = 0; i < lists.size(); ++i) printf("%s%s", i ? "," : "", lists[i].c_str());
  printf("],\"probe\":%s,\"unknown\":%s,\"version\":%s,\"health\":%s,\"log\":[", probe.body.c_str(), unknown.body.c_str(),
         api.Version().body.c_str(), api.Health().body.c_str());
  for (size_t i = 0; i < log.size(); ++i) printf("%s\"%s\"", i ? "," : "", jsonEscape(log[i]).c_str());
  printf("],\"metrics\":\"%s\"}\n", jsonEscape(m).c_str());
  return 0;
}
