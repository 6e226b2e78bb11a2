// host_selftest — host-side logic without a GPU: the device model mirror, grading dispatch
// with a scripted backend, the ListAndWatch feed, metrics text and the HTTP envelopes.
// Exits 0 and prints "host_selftest ok" when everything holds.  Run by tests/test_host.py.
#include <cassert>
#include <cstdio>
#include <thread>

#include "config.hpp"
#include "manager.hpp"
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
  std::set<std::string> bad, err, busy;
  int many_calls = 0, slice_calls = 0;
  int OtherComputeProcs(const std::string& u) override { return unknown_busy ? -1 : (busy.count(u) ? 2 : 0); }
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
      if (nomem.count(u)) { out->back().error = GP_ERR_NOMEM; out->back().healthy = 0; rc = GP_ERR_NOMEM; }
      if (err.count(u)) { out->back().error = GP_ERR_CUDA; out->back().healthy = 0; rc = GP_ERR_CUDA; if (e) *e = u + ": xid"; }
    }
    return rc;
  }
  // fabric script: flat 780 GB/s except the pairs in `slow_pairs`
  std::vector<std::pair<int, int>> slow_pairs;
  std::vector<std::string> fabric_uuids;
  int fabric_calls = 0, isolated_calls = 0, resets = 0;
  std::vector<std::string> trimmed;
  int unknown_busy = 0;  // OtherComputeProcs answers -1 for everything
  std::set<std::string> nomem;
  int ProbeFabric(int n, gp_fabric_result* out, std::vector<std::string>* uuids, std::string*) override {
    ++fabric_calls;
    gp_fabric_result f{};
    f.struct_size = sizeof(f); f.ngpus = (uint32_t)n; f.bytes = 1ull << 30; f.data_ok = 1; f.healthy = 1; f.spread = 1.0;
    f.worst_src = f.worst_dst = -1; f.worst_gbps = 780;
    for (int i = 0; i < n * n; ++i) { f.read_gbps[i] = 780; f.write_gbps[i] = 713; }
    for (auto pr : slow_pairs) {
      f.read_gbps[pr.first * n + pr.second] = f.read_gbps[pr.second * n + pr.first] = 90;
      f.write_gbps[pr.first * n + pr.second] = f.write_gbps[pr.second * n + pr.first] = 85;
      f.worst_src = pr.first; f.worst_dst = pr.second; f.worst_gbps = 85; f.spread = 780.0 / 85; f.failed = 1 | 2; f.healthy = 0;
    }
    f.allreduce_busbw_gbps = 830;
    *out = f;
    if (uuids) *uuids = fabric_uuids;
    return 0;
  }
  int ProbeIsolated(const std::string& u, uint32_t, uint64_t, gp_device_result* out, std::string*) override {
    ++isolated_calls;
    *out = mk(u, !bad.count(u));
    return 0;
  }
  void TrimDevice(const std::string& u) override { trimmed.push_back(u); }
  int ResetDevice(const std::string&) override { ++resets; return 0; }
  uint64_t last_slice_mem = 0;
  int ProbeSlice(const std::string& u, uint32_t, uint64_t total_mem, gp_device_result* out, std::string*) override {
    ++slice_calls;
    last_slice_mem = total_mem;
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
  CHECK(be.last_slice_mem == (180ull << 30));  // the slice's TotalMemory reaches the backend
  CHECK(GpuProbeBackend::SliceBytes(0, 23ull << 30) == ((23ull << 30) / 4 & ~((1ull << 20) - 1)));  // 1g.23gb: 5.75 GiB arrays
  CHECK(GpuProbeBackend::SliceBytes(1ull << 30, 180ull << 30) == (1ull << 30) && GpuProbeBackend::SliceBytes(0, 0) == (16ull << 30));
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

  // --- NewDeviceMap / strategies / error text (device/device_map.go:24-98) ---
  auto mkinfo = [](const std::string& u) { auto p = std::make_shared<FakeInfo>(); p->uuid = u; return p; };
  std::vector<GpuNode> nodes;
  for (int i = 0; i < 3; ++i) nodes.push_back({i, "NVIDIA B200", false, mkinfo("GPU-n" + std::to_string(i)), {}});
  nodes[2].migEnabled = true;
  nodes[2].migs = {{0, "1g.23gb", mkinfo("MIG-a")}, {1, "1g.23gb", mkinfo("MIG-b")}, {2, "7g.180gb", mkinfo("MIG-c")}};
  {
    auto [m1, e1] = NewDeviceMap(nodes, {Resource::New("*", "gpu")}, MigStrategyNone);
    CHECK(!e1 && m1.m["nvidia.com/gpu"].size() == 3);                       // none: MIG-enabled GPUs stay whole
    auto [m2, e2] = NewDeviceMap(nodes, {Resource::New("*", "gpu")}, MigStrategySingle);
    CHECK(!e2 && m2.m["nvidia.com/gpu"].size() == 2);                       // single: MIG-enabled GPU skipped
    auto [m3, e3] = NewDeviceMap(nodes, {Resource::New("1g.23gb", "mig-1g.23gb"), Resource::New("7g.180gb", "mig-7g.180gb")}, MigStrategyMixed);
    CHECK(!e3 && m3.m["nvidia.com/mig-1g.23gb"].size() == 2 && m3.m["nvidia.com/mig-7g.180gb"].GetByIndex("2:2"));
    auto [m4, e4] = NewDeviceMap(nodes, {Resource::New("1g.23gb", "mig-1g.23gb")}, MigStrategyMixed);
    CHECK(e4 && *e4 == "MIG profile '7g.180gb' does not match any resource patterns");
    auto [m5, e5] = NewDeviceMap(nodes, {Resource::New("GPU", "nvidia.com/gpu")}, MigStrategyNone);  // the reference's own pattern
    CHECK(e5 && *e5 == "GPU name 'NVIDIA B200' does not match any resource patterns");
    auto [m6, e6] = NewDeviceMap(nodes, {}, "weird");
    CHECK(e6 && *e6 == "invalid MIG strategy: weird");
    (void)m4; (void)m5; (void)m6;
  }
  // --- mig-minors parser (device/mig.go) ---
  {
    auto caps = GetMigCapabilityDevicePaths("config 1\nmonitor 2\ngpu0/gi1/access 12\ngpu0/gi1/ci0/access 13\ngarbage line\n");
    CHECK(caps.size() == 4 && caps["/proc/driver/nvidia/capabilities/gpu0/mig/gi1/ci0/access"] == "/dev/nvidia-caps/nvidia-cap13");
    CHECK(caps["/proc/driver/nvidia/capabilities/gpu0/mig/gi1/access"] == "/dev/nvidia-caps/nvidia-cap12" && caps["/proc/driver/nvidia/capabilities/mig/config"] == "/dev/nvidia-caps/nvidia-cap1");
  }
  // --- PluginManager: grade inside loadPlugins, Restart flag, Reprobe feeds ListAndWatch, ticker ---
  {
    FakeBackend fb;
    fb.bad = {"GPU-n1"};
    ProbeMetrics pm;
    PluginManager mgr(MigStrategySingle, {Resource::New("*", "gpu")}, [&] { return nodes; }, &fb, &pm, GP_PROBE_ALL_SINGLE);
    CHECK(!mgr.loadPlugins() && mgr.loads() == 1 && mgr.unhealthyAtLoad() == 1 && mgr.plugins().size() == 1);
    auto plug0 = mgr.plugins()[0];
    std::vector<std::vector<PluginDevice>> got;
    std::thread t([&] { plug0->ListAndWatch([&](const std::vector<PluginDevice>& l) { got.push_back(l); return true; }); });
    fb.bad = {"GPU-n0"};                        // n1 recovers, n0 breaks
    auto [graded, re] = mgr.Reprobe();
    CHECK(!re && graded.size() == 2);
    CHECK(mgr.Reprobe("GPU-zzz").second.value() == "unknown device GPU-zzz");
    plug0->Stop();
    t.join();
    CHECK(got.size() == 3);                     // first list + two transitions
    CHECK(health_of(got[0], "GPU-n1") == "Unhealthy" && health_of(got[2], "GPU-n1") == "Healthy" && health_of(got[2], "GPU-n0") == "Unhealthy");
    CHECK(pm.Render().find("gpu_probe_healthy{index=\"0\",uuid=\"GPU-n0\"} 0") != std::string::npos);
    mgr.Restart();
    CHECK(!mgr.poll() && mgr.loads() == 2 && !mgr.poll() && mgr.loads() == 2);   // flag consumed once
    const int before = mgr.reprobes();
    mgr.StartTicker(std::chrono::milliseconds(5));
    std::this_thread::sleep_for(std::chrono::milliseconds(80));
    mgr.StopTicker();
    CHECK(mgr.reprobes() >= before + 3);
    // the periodic path leaves a busy device alone; the on-demand path does not
    fb.busy = {"GPU-n0"};
    const int calls = fb.many_calls;
    auto [quiet, qe] = mgr.Reprobe("GPU-n0", /*skip_busy=*/true);
    CHECK(!qe && quiet.empty() && fb.many_calls == calls && mgr.skippedBusy() >= 1);
    auto [forced, fe2] = mgr.Reprobe("GPU-n0");
    CHECK(!fe2 && forced.size() == 1 && fb.many_calls == calls + 1);
    // with a reduced engine configured, a busy device gets the small probe instead of none
    FakeBackend small;
    mgr.SetReducedBackend(&small, GP_PROBE_COPY | GP_PROBE_PATTERN);
    auto [red, re3] = mgr.Reprobe("", /*skip_busy=*/true);
    CHECK(!re3 && red.size() == 2 && small.many_calls == 1 && mgr.reducedProbes() == 1 && fb.many_calls == calls + 2);
    PluginManager bad(MigStrategyNone, {Resource::New("GPU", "gpu")}, [&] { return nodes; }, &fb, &pm, 1);
    CHECK(bad.loadPlugins().value() == "GPU name 'NVIDIA B200' does not match any resource patterns");
  }
  // --- loadPlugins / restartPlugins never give the full probe to a device tenants are using ---
  {
    FakeBackend fb;
    ProbeMetrics pm;
    PluginManager mgr(MigStrategySingle, {Resource::New("*", "gpu")}, [&] { return nodes; }, &fb, &pm, GP_PROBE_ALL_SINGLE);
    mgr.EnableFabric(false);
    fb.bad = {"GPU-n1"};
    CHECK(!mgr.loadPlugins() && mgr.unhealthyAtLoad() == 1);      // idle node: both graded, n1 fails
    fb.bad.clear();
    fb.busy = {"GPU-n0", "GPU-n1"};                                 // pods arrive; kubelet restarts -> restartPlugins
    const int calls = fb.many_calls;
    CHECK(!mgr.restartPlugins() && fb.many_calls == calls);         // nobody was probed
    auto& g = mgr.devices().m["nvidia.com/gpu"];
    CHECK(g.GetByID("GPU-n0")->Health == std::string(Healthy));    // kept: last verdict healthy
    CHECK(g.GetByID("GPU-n1")->Health == std::string(Unhealthy));  // kept: last verdict unhealthy, not silently healed
    CHECK(mgr.skippedBusy() == 2 && mgr.unhealthyAtLoad() == 0);
    FakeBackend small;
    mgr.SetReducedBackend(&small, GP_PROBE_COPY | GP_PROBE_PATTERN);
    CHECK(!mgr.restartPlugins() && small.many_calls == 1 && fb.many_calls == calls && mgr.reducedProbes() == 2);
    // NVML missing: busy-ness unknown.  The full probe is tried; out of memory then means "in use", not "broken"
    FakeBackend blind;
    blind.unknown_busy = 1;
    blind.nomem = {"GPU-n0"};
    PluginManager m2(MigStrategySingle, {Resource::New("*", "gpu")}, [&] { return nodes; }, &blind, &pm, GP_PROBE_ALL_SINGLE);
    m2.EnableFabric(false);
    CHECK(!m2.loadPlugins() && m2.unhealthyAtLoad() == 0 && m2.notGraded() == 1);
    CHECK(m2.devices().m["nvidia.com/gpu"].GetByID("GPU-n0")->Health == std::string(Healthy));
    blind.unknown_busy = 0;                                         // with NVML answering "idle", NOMEM is a real failure
    PluginManager m3(MigStrategySingle, {Resource::New("*", "gpu")}, [&] { return nodes; }, &blind, &pm, GP_PROBE_ALL_SINGLE);
    m3.EnableFabric(false);
    CHECK(!m3.loadPlugins() && m3.unhealthyAtLoad() == 1);
  }
  // --- fabric: measured once per load, feeds the gauges and the allocator, steers around a degraded pair ---
  {
    std::vector<GpuNode> four;
    for (int i = 0; i < 4; ++i) four.push_back({i, "NVIDIA B200", false, mkinfo("GPU-f" + std::to_string(i)), {}});
    FakeBackend fb;
    fb.fabric_uuids = {"GPU-f0", "GPU-f1", "GPU-f2", "GPU-f3"};
    fb.slow_pairs = {{0, 1}};
    ProbeMetrics pm;
    PluginManager mgr(MigStrategyNone, {Resource::New("*", "gpu")}, [&] { return four; }, &fb, &pm, GP_PROBE_ALL_SINGLE);
    CHECK(!mgr.loadPlugins() && fb.fabric_calls == 1 && mgr.fabric().healthy == 0 && mgr.fabric().worst_src == 0);
    const std::string text = pm.Render();
    CHECK(text.find("gpu_probe_p2p_gbps{dir=\"read\",dst=\"1\",src=\"0\"} 90") != std::string::npos);
    CHECK(text.find("gpu_probe_p2p_gbps{dir=\"write\",dst=\"3\",src=\"2\"} 713") != std::string::npos);
    CHECK(text.find("gpu_probe_allreduce_busbw_gbps{ranks=\"4\"} 830") != std::string::npos);
    CHECK(text.find("gpu_probe_fabric_healthy{gpus=\"4\"} 0") != std::string::npos);
    bool logged = false;
    for (auto& l : mgr.log()) logged = logged || l.find("fabric graded unhealthy") != std::string::npos;
    CHECK(logged);
    auto [pick, pe] = mgr.PreferredAllocation("nvidia.com/gpu", fb.fabric_uuids, {"GPU-f0"}, 3);
    CHECK(!pe && pick.size() == 3 && pick[0] == "GPU-f0" && std::find(pick.begin(), pick.end(), "GPU-f1") == pick.end());
    auto [pick2, pe2] = mgr.PreferredAllocation("nvidia.com/gpu", fb.fabric_uuids, {}, 2);
    CHECK(!pe2 && !(std::count(pick2.begin(), pick2.end(), "GPU-f0") && std::count(pick2.begin(), pick2.end(), "GPU-f1")));
    CHECK(mgr.PreferredAllocation("nvidia.com/nope", {}, {}, 1).second.has_value());
    // busy GPUs: the fabric probe (it drives every GPU at once) is not run
    fb.busy = {"GPU-f2"};
    CHECK(!mgr.restartPlugins() && fb.fabric_calls == 1);
    // Allocate hands the kept probe arena back before the pod starts
    auto [env2, ae4] = mgr.plugins()[0]->Allocate({"GPU-f1", "GPU-f3"});
    CHECK(!ae4 && env2 == "NVIDIA_VISIBLE_DEVICES=GPU-f1,GPU-f3" && fb.trimmed == std::vector<std::string>({"GPU-f1", "GPU-f3"}));
  }
  // --- a context lost to a sticky fault: re-graded from a child process, then reset in-process ---
  {
    FakeBackend fb;
    ProbeMetrics pm;
    PluginManager mgr(MigStrategySingle, {Resource::New("*", "gpu")}, [&] { return nodes; }, &fb, &pm, GP_PROBE_ALL_SINGLE);
    mgr.EnableFabric(false);
    CHECK(!mgr.loadPlugins());
    auto plug0 = mgr.plugins()[0];
    std::vector<std::vector<PluginDevice>> got;
    std::thread t([&] { plug0->ListAndWatch([&](const std::vector<PluginDevice>& l) { got.push_back(l); return true; }); });
    auto lost = std::make_shared<gp_device_result>(FakeBackend::mk("GPU-n0", false));
    lost->error = GP_ERR_DEVICE_LOST;
    mgr.devices().m["nvidia.com/gpu"].GetByID("GPU-n0")->Probe = lost;   // what an Xid during a probe leaves behind
    plug0->Report("GPU-n0", false);
    const int many = fb.many_calls;
    auto [gr, ge] = mgr.Reprobe();
    CHECK(!ge && gr.size() == 2 && fb.isolated_calls == 1 && fb.resets == 1 && fb.many_calls == many + 1);  // n1 in-process, n0 isolated
    CHECK(mgr.isolatedProbes() == 1 && VerdictHealthy(*mgr.devices().m["nvidia.com/gpu"].GetByID("GPU-n0")));
    mgr.SetIsolation(true);                                            // policy: every periodic re-probe isolated
    mgr.Reprobe();
    CHECK(fb.isolated_calls == 3 && fb.many_calls == many + 1 && fb.resets == 1);
    plug0->Stop();
    t.join();
    CHECK(got.size() == 3 && health_of(got[1], "GPU-n0") == "Unhealthy" && health_of(got[2], "GPU-n0") == "Healthy");
  }
  // --- Start(): blocks instead of spinning, serves restarts, returns when stopped (the reference's loop does neither) ---
  {
    FakeBackend fb;
    ProbeMetrics pm;
    PluginManager mgr(MigStrategySingle, {Resource::New("*", "gpu")}, [&] { return nodes; }, &fb, &pm, GP_PROBE_ALL_SINGLE);
    mgr.EnableFabric(false);
    Error start_err = std::string("not returned");
    std::thread runner([&] { start_err = mgr.Start(); });
    for (int i = 0; i < 200 && !mgr.running(); ++i) std::this_thread::sleep_for(std::chrono::milliseconds(5));
    CHECK(mgr.running() && mgr.loads() == 1);
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
    CHECK(mgr.wakeups() == 0);                       // idle means asleep: no `default:` arm spinning a core
    mgr.Restart();
    for (int i = 0; i < 200 && mgr.loads() < 2; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(5));
    CHECK(mgr.loads() == 2 && mgr.wakeups() == 1);   // one wake-up per request, flag consumed once
    auto plug0 = mgr.plugins()[0];
    std::thread law([&] { plug0->ListAndWatch([](const std::vector<PluginDevice>&) { return true; }); });
    mgr.Stop();
    runner.join();                                   // Start() returns after stopping the plugins ...
    law.join();                                      // ... whose ListAndWatch streams have ended
    CHECK(!start_err && !mgr.running());
    PluginManager bad(MigStrategyNone, {Resource::New("GPU", "gpu")}, [&] { return nodes; }, &fb, &pm, 1);
    CHECK(bad.Start().value() == "GPU name 'NVIDIA B200' does not match any resource patterns");  // load errors come back, not a hang
  }
  // --- preferred allocation ---
  {
    Devices reps;
    for (const char* id : {"GPU-a::0", "GPU-a::1", "GPU-b::0", "GPU-b::1"}) { auto d = std::make_shared<Device>(); d->ID = id; reps.m[id] = d; }
    auto [alloc, ae2] = distributedAlloc(reps, {"GPU-a::0", "GPU-a::1", "GPU-b::0", "GPU-b::1"}, {}, 2);
    CHECK(!ae2 && alloc.size() == 2 && AnnotatedID{alloc[0]}.GetID() != AnnotatedID{alloc[1]}.GetID());  // spread over both GPUs
    CHECK(distributedAlloc(reps, {"GPU-a::0"}, {}, 2).second.value() == "not enough available devices to satisfy allocation");
    auto [withreq, ae3] = distributedAlloc(reps, {"GPU-a::0", "GPU-a::1", "GPU-b::0"}, {"GPU-a::0"}, 2);
    CHECK(!ae3 && withreq.size() == 2 && withreq[0] == "GPU-a::0" && withreq[1] != "GPU-a::0");  // required first; tie broken as the reference does
    std::vector<std::string> u4 = {"G0", "G1", "G2", "G3"};
    std::vector<double> bw(16, 780.0);
    bw[0 * 4 + 1] = bw[1 * 4 + 0] = 90.0;       // a degraded 0<->1 link
    auto [pick, pe] = alignedAllocFromMatrix(u4, bw, u4, {"G0"}, 3);
    CHECK(!pe && pick.size() == 3 && pick[0] == "G0" && std::find(pick.begin(), pick.end(), "G1") == pick.end());
    auto [flat, fe] = alignedAllocFromMatrix(u4, std::vector<double>(16, 780.0), u4, {}, 2);
    CHECK(!fe && flat == std::vector<std::string>({"G0", "G1"}));
    CHECK(alignedAllocFromMatrix(u4, bw, {"G0"}, {}, 2).second.has_value());
    // getPreferredAllocation picks the branch as plugin.go:248-254 does
    Devices whole;
    for (const auto& u : u4) { auto d = std::make_shared<Device>(); d->ID = u; d->Index = u.substr(1); d->Paths = {"/dev/nvidia" + u.substr(1)}; whole.m[u] = d; }
    auto [p1, pe1] = getPreferredAllocation(whole, P2PMatrix{u4, bw}, u4, {"G1"}, 2);
    CHECK(!pe1 && p1[0] == "G1" && p1[1] != "G0");                 // steers around the degraded 0<->1 link
    auto [p2, pe2] = getPreferredAllocation(whole, P2PMatrix{}, {"G3", "G2"}, {}, 1);
    CHECK(!pe2 && p2 == std::vector<std::string>({"G2"}));         // no matrix yet: id order
    auto [p3, pe3] = getPreferredAllocation(reps, P2PMatrix{u4, bw}, {"GPU-a::0", "GPU-a::1", "GPU-b::0", "GPU-b::1"}, {}, 2);
    CHECK(!pe3 && AnnotatedID{p3[0]}.GetID() != AnnotatedID{p3[1]}.GetID());  // annotated ids -> distributedAlloc
  }
  // --- config: the reference's own config.yml, byte for byte, plus optional probe keys ---
  {
    const char* ref_yml =
        "# web listen address\nwebListenAddress: \"0.0.0.0:9100\"\n\n# mig strategy\nmigStrategy: \"none\"\n\n"
        "# enable benchmark\nbenchmark: false\n\n# log configuration\nlog:\n    level: \"debug\"\n    fileDir: \"./logs\"\n";
    Config c = LoadConfig(ref_yml);
    CHECK(c.WebListenAddress == "0.0.0.0:9100" && c.MigStrategy == "none" && !c.Benchmark && c.Log.Level == "debug" && c.Log.FileDir == "./logs");
    CHECK(c.Probe.Enabled && c.Probe.HbmBytes == (16ull << 30) && c.Probe.WoPasses == 64 && c.Probe.IntervalSeconds == 0);
    Config d = LoadConfig("");
    CHECK(d.WebListenAddress == "9002" && d.MigStrategy == "none" && d.Log.Level == "debug");  // config.go:16-22
    Config e3 = LoadConfig(std::string(ref_yml) + "probe:\n  enabled: true\n  hbmBytes: 1073741824   # 1 GiB\n  minCopyGbps: 6000\n  intervalSeconds: 300\n");
    CHECK(e3.Probe.HbmBytes == (1ull << 30) && e3.Probe.MinCopyGbps == 6000 && e3.Probe.IntervalSeconds == 300 && e3.MigStrategy == "none");
    gp_config g = e3.ToGpConfig();
    CHECK(g.struct_size == sizeof(gp_config) && g.hbm_bytes == (1ull << 30) && g.min_copy_gbps == 6000 && g.gemm_output == 0);
    CHECK(LoadConfig("probe:\n  gemmOutput: bf16\n").ToGpConfig().gemm_output == 1);
    CHECK(LoadConfig("probe:\n  patternBytes: all\n").ToGpConfig().pattern_bytes == GP_PATTERN_ALL_FREE);
    Config f3 = LoadConfig("probe:\n  minP2PGbps: 600\n  maxP2PSpread: 1.2\n  minAllreduceBusbwGbps: 500\n  fabricBytes: 268435456\n  keepArena: true\n  isolate: true\n  fabric: false\n");
    gp_config fg = f3.ToGpConfig();
    CHECK(fg.min_p2p_gbps == 600 && fg.max_p2p_spread == 1.2 && fg.min_allreduce_busbw_gbps == 500 && fg.fabric_bytes == (256ull << 20) && fg.keep_arena == 1);
    Config k3 = LoadConfig("probe:\n  triadExpectSum: 0xb4e518db1999320e\n  triadExpectXor: 0xb4fb50e5795c70ba\n");
    CHECK(k3.ToGpConfig().triad_expect_sum == 0xb4e518db1999320eull && k3.ToGpConfig().triad_expect_xor == 0xb4fb50e5795c70baull);
    CHECK(f3.Probe.Isolate && !f3.Probe.Fabric && LoadConfig("").Probe.Fabric && LoadConfig("").ToGpConfig().keep_arena == 0);
  }
  printf("host_selftest ok\n");
  return 0;
}
