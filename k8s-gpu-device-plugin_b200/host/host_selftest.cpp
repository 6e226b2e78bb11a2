// host_selftest — host-side logic without a GPU: the device model mirror, grading dispatch
// with a scripted backend, the ListAndWatch feed, metrics text and the HTTP envelopes.
// Exits 0 and prints "host_selftest ok" when everything holds.  Run by tests/test_host.py.
#include <cassert>
#include <cstdio>
#include <thread>

#include "router.hpp"

using namespace plugin;

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

struct FakeInfo : deviceInfo {
  std::string uuid; bool fail_paths = false; bool numa = false;
  std::pair<std::string, Error> GetUUID() override { return {uuid, std::nullopt}; }
  std::pair<std::vector<std::string>, Error> GetPaths() override {
    if (fail_paths) return {{}, std::string("boom")};
    return {{"/dev/nvidia0"}, std::nullopt};
  }
  std::tuple<bool, int, Error> GetNumaNode() override { return {numa, 1, std::nullopt}; }
  std::pair<uint64_t, Error> GetTotalMemory() override { return {180ull << 30, std::nullopt}; }
  std::pair<std::string, Error> GetComputeCapability() override { return {"10.0", std::nullopt}; }
};

// scripted engine: `bad` uuids fail their verdict, `err` uuids make the call itself fail
struct FakeBackend : ProbeBackend {
  std::set<std::string> bad, err;
  int many_calls = 0, slice_calls = 0;
  static gp_device_result mk(const std::string& u, bool healthy) {
    gp_device_result r{};
    r.struct_size = sizeof(r);
    snprintf(r.uuid, sizeof(r.uuid), "%s", u.c_str());
    r.healthy = healthy; r.failed_mask = healthy ? 0 : GP_PROBE_PATTERN; r.probed_mask = GP_PROBE_ALL_SINGLE;
    r.copy_gbps_median = 6900; r.triad_gbps_median = 7100; r.fill_gbps_median = 7600; r.verify_gbps_median = 7500;
    r.pattern_passes = 66; r.pattern_mismatches = healthy ? 0 : 3; r.gemm.tflops_median = 1450; r.duration_s = 0.6;
    return r;
  }
  int ProbeMany(const std::vector<std::string>& uuids, uint32_t, std::vector<gp_device_result>* out, std::string* e) override {
    ++many_calls;
    out->clear();
    int rc = 0;
    for (auto& u : uuids) {
      out->push_back(mk(u, !bad.count(u)));
      if (err.count(u)) { out->back().error = GP_ERR_CUDA; out->back().healthy = 0; rc = GP_ERR_CUDA; if (e) *e = u + ": xid"; }
    }
    return rc;
  }
  int ProbeSlice(const std::string& u, uint32_t, gp_device_result* out, std::string*) override {
    ++slice_calls;
    *out = mk(u, !bad.count(u));
    return 0;
  }
};

int main() {
  // --- BuildDevice / Devices / AnnotatedID (device/devices.go) ---
  FakeInfo a; a.uuid = "GPU-aaaa"; a.numa = true;
  auto [dev, e] = BuildDevice("0", a);
  CHECK(!e && dev->ID == "GPU-aaaa" && dev->Index == "0" && dev->Health == std::string(Healthy));
  CHECK(dev->NumaNode && *dev->NumaNode == 1 && dev->ComputeCapability == "10.0" && !dev->IsMigDevice());
  FakeInfo bad; bad.uuid = "GPU-x"; bad.fail_paths = true;
  auto [d2, e2] = BuildDevice("1", bad);
  CHECK(!d2 && e2 && *e2 == "error getting device paths: boom");
  CHECK(AnnotatedID::New("GPU-a", 3).s == "GPU-a::3" && AnnotatedID{"GPU-a::3"}.Split().second == 3);
  CHECK(AnnotatedID{"GPU-a"}.GetID() == "GPU-a" && !AnnotatedID{"GPU-a"}.HasAnnotations());
  CHECK(wildCardToRegexp("NVIDIA*B200") == "NVIDIA.*B200" && wildCardToRegexp("a.b*") == "a\\.b.*");
  CHECK(matchResourcePattern("*", "NVIDIA B200") && matchResourcePattern("*B200", "NVIDIA B200") && !matchResourcePattern("A100*", "NVIDIA B200"));

  // --- DeviceMap + GradeDevices: GPUs together, slices via ProbeSlice, errors => Unhealthy ---
  DeviceMap dm;
  FakeInfo g0, g1, g2, s0, s1;
  g0.uuid = "GPU-0"; g1.uuid = "GPU-1"; g2.uuid = "GPU-2"; s0.uuid = "MIG-0"; s1.uuid = "MIG-1";
  CHECK(!dm.setEntry("nvidia.com/gpu", "0", g0) && !dm.setEntry("nvidia.com/gpu", "1", g1) && !dm.setEntry("nvidia.com/gpu", "2", g2));
  CHECK(!dm.setEntry("nvidia.com/mig-1g.23gb", "3:0", s0) && !dm.setEntry("nvidia.com/mig-1g.23gb", "3:1", s1));
  CHECK(dm.setEntry("nvidia.com/gpu", "9", bad).value() == "error building Device: error getting device paths: boom");
  FakeBackend be;
  be.bad = {"GPU-1", "MIG-1"};
  be.err = {"GPU-2"};
  std::vector<std::string> log;
  CHECK(GradeDevices(dm, be, GP_PROBE_ALL_SINGLE, &log) == 3);
  CHECK(be.many_calls == 1 && be.slice_calls == 2 && log.size() == 3);
  auto& gpus = dm.m["nvidia.com/gpu"];
  CHECK(gpus.GetByID("GPU-0")->Health == std::string(Healthy) && gpus.GetByID("GPU-1")->Health == std::string(Unhealthy));
  CHECK(gpus.GetByID("GPU-2")->Health == std::string(Unhealthy) && gpus.GetByID("GPU-2")->Probe->error == GP_ERR_CUDA);
  CHECK(dm.m["nvidia.com/mig-1g.23gb"].GetByIndex("3:1")->Health == std::string(Unhealthy));
  CHECK(dm.m["nvidia.com/mig-1g.23gb"].GetByIndex("3:0")->Health == std::string(Healthy));
  CHECK(gpus.Contains({"GPU-0", "GPU-1"}) && !gpus.Contains({"GPU-0", "nope"}));
  CHECK(gpus.Subset({"GPU-0", "nope"}).size() == 1 && gpus.Difference(gpus.Subset({"GPU-0"})).size() == 2);
  CHECK(gpus.GetPluginDevices().size() == 3 && gpus.AlignedAllocationSupported());
  CHECK(!dm.m["nvidia.com/mig-1g.23gb"].AlignedAllocationSupported());

  // --- ListAndWatch feed: first list, one resend per real transition, both directions ---
  NvidiaDevicePlugin plug("nvidia.com/gpu", gpus);
  std::vector<std::vector<PluginDevice>> sent;
  std::thread law([&] { plug.ListAndWatch([&](const std::vector<PluginDevice>& l) { sent.push_back(l); return true; }); });
  plug.Report("GPU-0", false);   // transition -> resend
  plug.Report("GPU-0", false);   // same state -> nothing
  plug.Report("GPU-1", true);    // recovery (the reference has no way back) -> resend
  plug.Report("nope", false);    // unknown id -> ignored
  plug.Stop();
  law.join();
  CHECK(sent.size() == 3);
  auto health_of = [](const std::vector<PluginDevice>& l, const std::string& id) { for (auto& d : l) if (d.ID == id) return d.Health; return std::string("?"); };
  CHECK(health_of(sent[0], "GPU-0") == "Healthy" && health_of(sent[1], "GPU-0") == "Unhealthy" && health_of(sent[2], "GPU-1") == "Healthy");
  auto [env, ae] = plug.Allocate({"GPU-0", "GPU-2"});
  CHECK(!ae && env == "NVIDIA_VISIBLE_DEVICES=GPU-0,GPU-2");
  CHECK(plug.Allocate({"GPU-7"}).second.value() == "invalid allocation request for nvidia.com/gpu");

  // --- metrics + router ---
  ProbeMetrics metrics;
  for (auto& d : dm.All()) metrics.Update(*d);
  metrics.SetP2P(0, 1, "read", 780.0);
  metrics.SetAllReduce(8, 725.0);
  const std::string text = metrics.Render();
  CHECK(text.find("# TYPE gpu_probe_hbm_copy_gbps gauge") != std::string::npos);
  CHECK(text.find("gpu_probe_healthy{index=\"1\",uuid=\"GPU-1\"} 0") != std::string::npos);
  CHECK(text.find("gpu_probe_pattern_mismatches{index=\"3:1\",uuid=\"MIG-1\"} 3") != std::string::npos);
  CHECK(text.find("gpu_probe_p2p_gbps{dir=\"read\",dst=\"1\",src=\"0\"} 780") != std::string::npos);
  int reports = 0, restarts = 0;
  API api(&dm, &be, &metrics, GP_PROBE_ALL_SINGLE, [&](const std::string&, bool) { ++reports; }, [&] { ++restarts; });
  CHECK(api.Version().body == "{\"code\":0,\"data\":\"version : 0.0.1\",\"msg\":\"success\"}");
  CHECK(api.Health().body == "{\"code\":0,\"data\":\"ok\",\"msg\":\"success\"}");
  CHECK(api.Restart().body == api.Health().body && restarts == 1);
  be.bad.clear(); be.err.clear();
  const Response one = api.Probe("GPU-2");
  CHECK(one.status == 200 && one.body.find("\"code\":0") == 1 && one.body.find("\"uuid\":\"GPU-2\"") != std::string::npos && reports == 1);
  CHECK(gpus.GetByID("GPU-2")->Probe->healthy == 1);
  CHECK(gpus.GetByID("GPU-2")->Health == std::string(Unhealthy));  // re-probe never writes Health itself: Report() does
  const Response all = api.Probe();
  CHECK(reports == 1 + 5 && all.body.find("MIG-0") != std::string::npos);
  CHECK(api.Probe("GPU-none").body == "{\"code\":404,\"data\":null,\"msg\":\"unknown device GPU-none\"}");
  CHECK(Failed(7, "a\"b").find("a\\\"b") != std::string::npos);
  printf("host_selftest ok\n");
  return 0;
}
