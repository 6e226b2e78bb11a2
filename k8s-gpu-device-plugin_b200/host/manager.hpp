// manager.hpp — the rest of the reference's host path around the probe:
//   NewDeviceMap / build / buildGPUDeviceMap / buildMigDeviceMap   <- device/device_map.go:24-98
//   GetMigCapabilityDevicePaths (line parser)                       <- device/mig.go:20-81
//   PluginManager loadPlugins / restartPlugins / Restart            <- plugin/manager.go:102-194
//   distributedAlloc                                                <- plugin/plugin.go:284-326
//   preferred allocation from the MEASURED P2P matrix (replaces the NVML link query of
//   alignedAlloc, plugin/plugin.go:256-282, whose nvmllib is never set — SURVEY.md §8f.3)
// What is new: grading inside loadPlugins (before anything is advertised), Reprobe, a ticker.
// The kubelet socket, gRPC server and fsnotify watcher are transport and are not mirrored.
#pragma once
#include <algorithm>
#include <cstdio>
#include <sstream>

#include "metrics.hpp"
#include "probe.hpp"

namespace plugin {

// resource/resource.go:15-40
inline const char* const MigStrategyNone = "none";
inline const char* const MigStrategySingle = "single";
inline const char* const MigStrategyMixed = "mixed";
struct Resource {
  std::string Pattern, Name;
  static Resource New(const std::string& pattern, std::string name) {  // resource.go:31-40
    if (name.rfind("nvidia.com/", 0) != 0) name = "nvidia.com/" + name;
    return {pattern, name};
  }
};

// What go-nvlib's VisitDevices / VisitMigDevices hand to the builder (device_map.go:50,80)
struct MigNode {
  int j;
  std::string profile;  // MigProfile.String(), e.g. "1g.23gb"
  std::shared_ptr<deviceInfo> info;
};
struct GpuNode {
  int i;
  std::string name;  // gpu.GetName()
  bool migEnabled = false;
  std::shared_ptr<deviceInfo> info;
  std::vector<MigNode> migs;
};

// device_map.go:24-98
inline std::pair<DeviceMap, Error> NewDeviceMap(const std::vector<GpuNode>& gpus, const std::vector<Resource>& resources,
                                                const std::string& migStrategy) {
  DeviceMap devices;
  if (migStrategy == MigStrategyNone || migStrategy == MigStrategySingle) {  // buildGPUDeviceMap :48-75
    for (const auto& gpu : gpus) {
      if (gpu.migEnabled && migStrategy != MigStrategyNone) continue;
      bool matched = false;
      for (const auto& r : resources) {
        if (matchResourcePattern(r.Pattern, gpu.name)) {
          if (auto e = devices.setEntry(r.Name, std::to_string(gpu.i), *gpu.info)) return {devices, e};
          matched = true;
          break;
        }
      }
      if (!matched) return {devices, "GPU name '" + gpu.name + "' does not match any resource patterns"};
    }
    return {devices, std::nullopt};
  }
  if (migStrategy == MigStrategyMixed) {  // buildMigDeviceMap :78-98
    for (const auto& gpu : gpus)
      for (const auto& mig : gpu.migs) {
        bool matched = false;
        for (const auto& r : resources) {
          if (matchResourcePattern(r.Pattern, mig.profile)) {
            if (auto e = devices.setEntry(r.Name, std::to_string(gpu.i) + ":" + std::to_string(mig.j), *mig.info)) return {devices, e};
            matched = true;
            break;
          }
        }
        if (!matched) return {devices, "MIG profile '" + mig.profile + "' does not match any resource patterns"};
      }
    return {devices, std::nullopt};
  }
  return {devices, "invalid MIG strategy: " + migStrategy};
}

// device/mig.go:38-66: one line of /proc/driver/nvidia-caps/mig-minors -> (capability path, minor)
inline std::optional<std::pair<std::string, int>> parseMigMinorsLine(const std::string& line) {
  int gpu, gi, ci, minor;
  const std::string caps = "/proc/driver/nvidia/capabilities";
  if (sscanf(line.c_str(), "gpu%d/gi%d/ci%d/access %d", &gpu, &gi, &ci, &minor) == 4)
    return {{caps + "/gpu" + std::to_string(gpu) + "/mig/gi" + std::to_string(gi) + "/ci" + std::to_string(ci) + "/access", minor}};
  if (sscanf(line.c_str(), "gpu%d/gi%d/access %d", &gpu, &gi, &minor) == 3)
    return {{caps + "/gpu" + std::to_string(gpu) + "/mig/gi" + std::to_string(gi) + "/access", minor}};
  if (sscanf(line.c_str(), "config %d", &minor) == 1) return {{caps + "/mig/config", minor}};
  if (sscanf(line.c_str(), "monitor %d", &minor) == 1) return {{caps + "/mig/monitor", minor}};
  return std::nullopt;  // "unparsable line": skipped, as the reference logs and continues
}
inline std::map<std::string, std::string> GetMigCapabilityDevicePaths(const std::string& minorsText) {
  std::map<std::string, std::string> out;
  std::istringstream in(minorsText);
  std::string line;
  while (std::getline(in, line))
    if (auto p = parseMigMinorsLine(line)) out[p->first] = "/dev/nvidia-caps/nvidia-cap" + std::to_string(p->second);
  return out;
}

// plugin/plugin.go:284-326 — spread an allocation over replicated GPUs, least-used first
inline std::pair<std::vector<std::string>, Error> distributedAlloc(const Devices& all, const std::vector<std::string>& available,
                                                                   const std::vector<std::string>& required, int size) {
  std::vector<std::string> candidates = all.Subset(available).Difference(all.Subset(required)).GetIDs();
  const int needed = size - (int)required.size();
  if ((int)candidates.size() < needed) return {{}, "not enough available devices to satisfy allocation"};
  struct Cnt { int total = 0, available = 0; };
  std::map<std::string, Cnt> replicas;
  for (const auto& c : candidates) replicas[AnnotatedID{c}.GetID()].available++;
  for (const auto& kv : all.m) {
    const std::string id = AnnotatedID{kv.first}.GetID();
    if (replicas.count(id)) replicas[id].total++;
  }
  std::vector<std::string> devices;
  for (int i = 0; i < needed; ++i) {
    std::stable_sort(candidates.begin(), candidates.end(), [&](const std::string& a, const std::string& b) {
      const Cnt& ca = replicas[AnnotatedID{a}.GetID()];
      const Cnt& cb = replicas[AnnotatedID{b}.GetID()];
      return ca.total - ca.available < cb.total - cb.available;
    });
    replicas[AnnotatedID{candidates[0]}.GetID()].available--;
    devices.push_back(candidates[0]);
    candidates.erase(candidates.begin());
  }
  std::vector<std::string> out(required);
  out.insert(out.end(), devices.begin(), devices.end());
  return {out, std::nullopt};
}

// Preferred allocation from the measured P2P matrix: greedily grow the required set with the
// candidate whose WORST measured link to the set is best (ties: lowest id).  On a healthy
// NVSwitch box the matrix is flat and this is id order; with a degraded link it steers around it.
inline std::pair<std::vector<std::string>, Error> alignedAllocFromMatrix(
    const std::vector<std::string>& uuidByOrdinal, const std::vector<double>& gbps /*n*n, row=src*/,
    const std::vector<std::string>& available, const std::vector<std::string>& required, int size) {
  const int n = (int)uuidByOrdinal.size();
  auto ord = [&](const std::string& id) { for (int i = 0; i < n; ++i) if (uuidByOrdinal[(size_t)i] == id) return i; return -1; };
  std::vector<std::string> chosen(required);
  std::vector<std::string> cand;
  for (const auto& a : available)
    if (std::find(required.begin(), required.end(), a) == required.end()) cand.push_back(a);
  std::sort(cand.begin(), cand.end());
  if ((int)(chosen.size() + cand.size()) < size) return {{}, "not enough available devices to satisfy allocation"};
  auto link = [&](const std::string& a, const std::string& b) {
    const int i = ord(a), j = ord(b);
    if (i < 0 || j < 0 || (int)gbps.size() < n * n) return 0.0;
    return std::min(gbps[(size_t)(i * n + j)], gbps[(size_t)(j * n + i)]);
  };
  while ((int)chosen.size() < size) {
    int best = -1;
    double best_score = -1;
    for (int c = 0; c < (int)cand.size(); ++c) {
      double worst = 1e30;
      for (const auto& s : chosen) worst = std::min(worst, link(cand[(size_t)c], s));
      if (chosen.empty()) {  // seed with the device whose weakest link overall is strongest
        worst = 1e30;
        for (const auto& o : cand) if (o != cand[(size_t)c]) worst = std::min(worst, link(cand[(size_t)c], o));
      }
      if (worst > best_score) { best_score = worst; best = c; }
    }
    chosen.push_back(cand[(size_t)best]);
    cand.erase(cand.begin() + best);
  }
  return {chosen, std::nullopt};
}

// plugin/plugin.go:192-207,248-254: GetPreferredAllocation -> getPreferredAllocation.  Aligned
// allocation when every device supports it and no id is annotated (a replica), else spread over
// replicas.  The aligned branch ranks by the measured P2P matrix instead of go-gpuallocator's NVML
// link query; with no matrix measured yet it degrades to id order (a flat matrix).
struct P2PMatrix {
  std::vector<std::string> uuidByOrdinal;
  std::vector<double> gbps;  // n*n, row = src
};
inline std::pair<std::vector<std::string>, Error> getPreferredAllocation(const Devices& devices, const P2PMatrix& mx,
                                                                         const std::vector<std::string>& available,
                                                                         const std::vector<std::string>& mustInclude, int size) {
  bool annotated = false;
  for (const auto& a : available) annotated = annotated || AnnotatedID{a}.HasAnnotations();
  if (devices.AlignedAllocationSupported() && !annotated) {
    if (mx.uuidByOrdinal.empty()) {
      P2PMatrix flat;
      flat.uuidByOrdinal = available;
      std::sort(flat.uuidByOrdinal.begin(), flat.uuidByOrdinal.end());
      flat.gbps.assign(flat.uuidByOrdinal.size() * flat.uuidByOrdinal.size(), 1.0);
      return alignedAllocFromMatrix(flat.uuidByOrdinal, flat.gbps, available, mustInclude, size);
    }
    return alignedAllocFromMatrix(mx.uuidByOrdinal, mx.gbps, available, mustInclude, size);
  }
  return distributedAlloc(devices, available, mustInclude, size);
}

// plugin/manager.go:20-194 without the sockets: owns the device map and one plugin per resource.
class PluginManager {
 public:
  using Enumerate = std::function<std::vector<GpuNode>()>;
  PluginManager(std::string migStrategy, std::vector<Resource> resources, Enumerate enumerate, ProbeBackend* be,
                ProbeMetrics* metrics, uint32_t mask)
      : migStrategy_(std::move(migStrategy)), resources_(std::move(resources)), enumerate_(std::move(enumerate)),
        be_(be), metrics_(metrics), mask_(mask) {}
  ~PluginManager() { StopTicker(); }

  // manager.go:156-174, plus the grade: kubelet never sees an ungraded device.
  //
  // restartPlugins runs this again on every kubelet socket re-creation and every GET /restart
  // (manager.go:177-194), i.e. on nodes whose GPUs are running pods.  A device somebody else is
  // computing on is therefore never given the full probe here (3 x 16 GiB and every SM would either
  // fail to allocate or trample the tenant): it gets the reduced probe when a reduced backend is set,
  // else it keeps the verdict of its last grade (Healthy if it never had one).  When busy-ness cannot
  // be told (no NVML) the full probe is tried, but running out of memory then means "in use, not
  // graded", not "unhealthy".
  Error loadPlugins() {
    auto [dm, err] = NewDeviceMap(enumerate_(), resources_, migStrategy_);
    if (err) return err;
    std::lock_guard<std::mutex> g(mu_);
    devices_ = std::move(dm);
    DeviceMap idle, busy;
    std::map<std::string, int> procs;
    int n_busy = 0;
    for (auto& kv : devices_.m)
      for (auto& dv : kv.second.m) {
        const std::string u = dv.second->GetUUID();
        if (!procs.count(u)) procs[u] = be_->OtherComputeProcs(u);
        if (procs[u] > 0) { busy.m[kv.first].m[dv.first] = dv.second; ++n_busy; }
        else idle.m[kv.first].m[dv.first] = dv.second;
      }
    unhealthy_at_load_ = 0;
    if (!idle.m.empty()) unhealthy_at_load_ += GradeDevices(idle, *be_, mask_, &log_);
    for (auto& d : idle.All())  // unknown busy-ness + out of memory = a tenant holds the memory: not graded
      if (procs[d->GetUUID()] < 0 && d->Probe && d->Probe->error == GP_ERR_NOMEM) { KeepLastVerdict(*d); --unhealthy_at_load_; ++not_graded_; }
    if (!busy.m.empty()) {
      if (reduced_be_) { unhealthy_at_load_ += GradeDevices(busy, *reduced_be_, reduced_mask_, &log_); reduced_probes_ += n_busy; }
      else for (auto& d : busy.All()) { KeepLastVerdict(*d); ++skipped_busy_; }
    }
    for (auto& d : devices_.All()) { metrics_->Update(*d); if (d->Probe) last_[d->GetUUID()] = d->Probe; }
    // the fabric is graded once per load, with every full GPU idle and healthy (it drives all of them at once)
    if (fabric_enabled_ && n_busy == 0) GradeFabric();
    plugins_.clear();
    for (auto& kv : devices_.m) {
      auto pl = std::make_shared<NvidiaDevicePlugin>(kv.first, kv.second);
      pl->SetOnAllocate([this](const std::vector<std::string>& ids) { for (auto& id : ids) be_->TrimDevice(AnnotatedID{id}.GetID()); });
      plugins_.push_back(pl);
    }
    ++loads_;
    return std::nullopt;
  }
  // plugin/plugin.go:192-207 GetPreferredAllocation for resource `name`, ranking by the measured matrix
  std::pair<std::vector<std::string>, Error> PreferredAllocation(const std::string& name, const std::vector<std::string>& available,
                                                                 const std::vector<std::string>& mustInclude, int size) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = devices_.m.find(name);
    if (it == devices_.m.end()) return {{}, "unknown resource " + name};
    return getPreferredAllocation(it->second, fabric_matrix_, available, mustInclude, size);
  }
  void EnableFabric(bool on) { fabric_enabled_ = on; }
  const gp_fabric_result& fabric() const { return fabric_; }
  const P2PMatrix& fabricMatrix() const { return fabric_matrix_; }
  // plugin/manager.go:56-99 Start(): load, serve restart requests until told to stop, stop the plugins, RETURN.
  // The reference's loop has a `default:` arm, so it busy-spins a core (manager.go:89-97), reads an unsynchronised
  // `restart` flag (manager.go:94 vs :109) and does not return after ctx.Done().  This is the minimal fix the Go
  // side needs as well: block until there is something to do, atomic flag, return at the end.
  Error Start() {
    if (auto e = loadPlugins()) return e;
    std::unique_lock<std::mutex> lk(run_mu_);
    running_ = true;
    for (;;) {
      run_cv_.wait(lk, [this] { return stop_requested_ || restart_.load(); });
      ++wakeups_;
      if (stop_requested_) break;
      restart_ = false;
      lk.unlock();
      if (auto e = restartPlugins()) { lk.lock(); running_ = false; return e; }
      lk.lock();
    }
    lk.unlock();
    { std::lock_guard<std::mutex> g(mu_); for (auto& p : plugins_) p->Stop(); }
    lk.lock();
    running_ = false;
    return std::nullopt;
  }
  void Stop() {
    { std::lock_guard<std::mutex> g(run_mu_); stop_requested_ = true; }
    run_cv_.notify_all();
  }
  bool running() { std::lock_guard<std::mutex> g(run_mu_); return running_; }
  int wakeups() const { return wakeups_; }
  // manager.go:108-110 sets a flag that the Start loop consumes; poll() is the same thing for callers without a loop
  void Restart() {
    { std::lock_guard<std::mutex> g(run_mu_); restart_ = true; }
    run_cv_.notify_all();
  }
  Error poll() {
    if (!restart_.exchange(false)) return std::nullopt;
    return restartPlugins();
  }
  Error restartPlugins() {  // manager.go:177-194
    { std::lock_guard<std::mutex> g(mu_); for (auto& p : plugins_) p->Stop(); plugins_.clear(); devices_ = DeviceMap(); }
    return loadPlugins();
  }
  // A second, small-footprint engine (small arrays, few passes) for devices tenants are using: with
  // it the periodic path gives busy devices the reduced probe instead of leaving them ungraded.
  void SetReducedBackend(ProbeBackend* reduced, uint32_t reduced_mask) { reduced_be_ = reduced; reduced_mask_ = reduced_mask; }

  // re-grade (all, or one uuid) while streams are open: transitions go through Report
  // skip_busy: devices on which other processes compute are not given the full probe; they get the
  // reduced one if a reduced backend is set, else keep their last verdict
  std::pair<std::vector<std::shared_ptr<Device>>, Error> Reprobe(const std::string& uuid = "", bool skip_busy = false) {
    std::lock_guard<std::mutex> g(mu_);
    DeviceMap scope, busy;
    bool known = false;
    for (auto& kv : devices_.m)
      for (auto& dv : kv.second.m)
        if (uuid.empty() || dv.second->GetUUID() == uuid) {
          known = true;
          if (skip_busy && be_->OtherComputeProcs(dv.second->GetUUID()) > 0) {  // -1 (unknown) counts as idle here: the arena request then decides
            if (reduced_be_) { busy.m[kv.first].m[dv.first] = dv.second; ++reduced_probes_; }
            else ++skipped_busy_;
            continue;
          }
          scope.m[kv.first].m[dv.first] = dv.second;
        }
    if (!known) return {{}, "unknown device " + uuid};
    if (scope.m.empty() && busy.m.empty()) return {{}, std::nullopt};
    // a device whose in-process context was lost to a sticky fault (or every device, with isolation on) is
    // graded from a fresh child process: the fault cannot reach this process, and a lost context cannot
    // keep a repaired GPU Unhealthy for ever.  If the child finds the GPU healthy, the in-process context
    // is reset so the fast path works again.
    DeviceMap inproc, isolated;
    for (auto& kv : scope.m)
      for (auto& dv : kv.second.m) {
        const bool lost = dv.second->Probe && dv.second->Probe->error == GP_ERR_DEVICE_LOST;
        if (!dv.second->IsMigDevice() && (isolate_ || lost)) isolated.m[kv.first].m[dv.first] = dv.second;
        else inproc.m[kv.first].m[dv.first] = dv.second;
      }
    if (!inproc.m.empty()) GradeDevices(inproc, *be_, mask_, &log_, /*apply_health=*/false);
    if (!isolated.m.empty()) GradeIsolated(isolated);
    if (!busy.m.empty()) GradeDevices(busy, *reduced_be_, reduced_mask_, &log_, /*apply_health=*/false);
    auto all = scope.All();
    for (auto& d : busy.All()) all.push_back(d);
    for (auto& d : all) {
      metrics_->Update(*d);
      if (d->Probe) last_[d->GetUUID()] = d->Probe;
      for (auto& p : plugins_) if (p->devices().GetByID(d->ID)) p->Report(d->ID, VerdictHealthy(*d));
    }
    ++reprobes_;
    return {all, std::nullopt};
  }
  // every full-GPU re-probe through a child process (slower: context creation and a cold arena per call)
  void SetIsolation(bool on) { isolate_ = on; }
  int isolatedProbes() const { return isolated_probes_; }
  int notGraded() const { return not_graded_; }

  void StartTicker(std::chrono::milliseconds interval) {
    StopTicker();
    ticking_ = true;
    ticker_ = std::thread([this, interval] {
      std::unique_lock<std::mutex> lk(tick_mu_);
      while (ticking_) {
        if (tick_cv_.wait_for(lk, interval, [this] { return !ticking_; })) break;
        lk.unlock();
        Reprobe("", /*skip_busy=*/true);
        lk.lock();
      }
    });
  }
  void StopTicker() {
    { std::lock_guard<std::mutex> g(tick_mu_); ticking_ = false; }
    tick_cv_.notify_all();
    if (ticker_.joinable()) ticker_.join();
  }
  std::vector<std::shared_ptr<NvidiaDevicePlugin>> plugins() { std::lock_guard<std::mutex> g(mu_); return plugins_; }
  DeviceMap& devices() { return devices_; }
  int loads() const { return loads_; }
  int reprobes() const { return reprobes_; }
  int skippedBusy() const { return skipped_busy_; }
  int reducedProbes() const { return reduced_probes_; }
  int unhealthyAtLoad() const { return unhealthy_at_load_; }
  std::vector<std::string> log() { std::lock_guard<std::mutex> g(mu_); return log_; }

 private:
  // not graded this time: the device keeps the verdict of its last grade (Healthy if it never had one)
  void KeepLastVerdict(Device& d) {
    auto it = last_.find(d.GetUUID());
    if (it != last_.end()) { d.Probe = it->second; d.Health = VerdictHealthy(d) ? Healthy : Unhealthy; }
    else { d.Probe = nullptr; d.Health = Healthy; }
  }
  void GradeIsolated(DeviceMap& dm) {
    auto devs = dm.All();
    std::vector<gp_device_result> res(devs.size());
    std::vector<int> rcs(devs.size(), 0);
    std::vector<std::string> errs(devs.size());
    std::vector<std::thread> th;
    for (size_t i = 0; i < devs.size(); ++i)
      th.emplace_back([&, i] { rcs[i] = be_->ProbeIsolated(devs[i]->GetUUID(), mask_, devs[i]->TotalMemory, &res[i], &errs[i]); });
    for (auto& t : th) t.join();
    for (size_t i = 0; i < devs.size(); ++i) {
      ++isolated_probes_;
      const bool ok = rcs[i] == GP_OK && res[i].error == 0 && res[i].healthy == 1;
      auto pr = std::make_shared<gp_device_result>(res[i]);
      if (!ok) { pr->healthy = 0; if (pr->error == 0) pr->error = rcs[i] ? rcs[i] : GP_ERR_CUDA; }
      const bool was_lost = devs[i]->Probe && devs[i]->Probe->error == GP_ERR_DEVICE_LOST;
      devs[i]->Probe = pr;
      if (!ok) log_.push_back("'" + devs[i]->ID + "' graded unhealthy from a child process: rc=" + std::to_string(rcs[i]) + (errs[i].empty() ? "" : " (" + errs[i] + ")"));
      else if (was_lost) {
        const int rr = be_->ResetDevice(devs[i]->GetUUID());
        log_.push_back("'" + devs[i]->ID + "' healthy again from a fresh process; in-process context reset rc=" + std::to_string(rr));
      }
    }
  }
  // K4 once per load: matrix for the allocator, gauges, verdict in the log
  void GradeFabric() {
    std::vector<std::string> full;
    for (auto& d : devices_.All())
      if (!d->IsMigDevice() && !AnnotatedID{d->ID}.HasAnnotations()) full.push_back(d->GetUUID());
    std::sort(full.begin(), full.end());
    full.erase(std::unique(full.begin(), full.end()), full.end());
    if (full.size() < 2 || full.size() > GP_FABRIC_MAX_GPUS) return;
    std::vector<std::string> uuids;
    std::string err;
    gp_fabric_result f{};
    const int rc = be_->ProbeFabric((int)full.size(), &f, &uuids, &err);
    if (rc != GP_OK) { log_.push_back("fabric probe failed: rc=" + std::to_string(rc) + (err.empty() ? "" : " (" + err + ")")); return; }
    fabric_ = f;
    const int n = (int)f.ngpus;
    fabric_matrix_.uuidByOrdinal = uuids;
    fabric_matrix_.gbps.assign((size_t)n * n, 0.0);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) {
        // what a pair can sustain in the slower of its two modes; the allocator takes min(i->j, j->i) of this
        fabric_matrix_.gbps[(size_t)(i * n + j)] = std::min(f.read_gbps[i * n + j], f.write_gbps[i * n + j]);
        if (i != j) { metrics_->SetP2P(i, j, "read", f.read_gbps[i * n + j]); metrics_->SetP2P(i, j, "write", f.write_gbps[i * n + j]); }
      }
    if (f.allreduce_error == 0) metrics_->SetAllReduce(n, f.allreduce_busbw_gbps);
    metrics_->Set("gpu_probe_fabric_healthy", "1 if the P2P matrix and the all-reduce met their floors and every transfer carried the right bytes", "gpus=\"" + std::to_string(n) + "\"", f.healthy);
    metrics_->Set("gpu_probe_fabric_spread", "max/min over the off-diagonal P2P pairs (1.0 = flat)", "gpus=\"" + std::to_string(n) + "\"", f.spread);
    if (!f.healthy)
      log_.push_back("fabric graded unhealthy: failed=" + std::to_string(f.failed) + " worst pair " + std::to_string(f.worst_src) + "->" +
                     std::to_string(f.worst_dst) + " at " + std::to_string(f.worst_gbps) + " GB/s, spread " + std::to_string(f.spread));
  }

  std::string migStrategy_;
  std::vector<Resource> resources_;
  Enumerate enumerate_;
  ProbeBackend* be_;
  ProbeMetrics* metrics_;
  uint32_t mask_;
  std::mutex mu_;
  DeviceMap devices_;
  std::vector<std::shared_ptr<NvidiaDevicePlugin>> plugins_;
  std::vector<std::string> log_;
  std::atomic<bool> restart_{false};
  std::atomic<int> loads_{0}, reprobes_{0}, skipped_busy_{0}, reduced_probes_{0};
  ProbeBackend* reduced_be_ = nullptr;
  uint32_t reduced_mask_ = GP_PROBE_COPY | GP_PROBE_PATTERN;
  int unhealthy_at_load_ = 0;
  std::map<std::string, std::shared_ptr<gp_device_result>> last_;  // by uuid, survives restartPlugins
  bool fabric_enabled_ = true, isolate_ = false;
  gp_fabric_result fabric_{};
  P2PMatrix fabric_matrix_;
  std::atomic<int> isolated_probes_{0}, not_graded_{0};
  std::mutex run_mu_;
  std::condition_variable run_cv_;
  bool stop_requested_ = false, running_ = false;
  std::atomic<int> wakeups_{0};
  std::thread ticker_;
  std::mutex tick_mu_;
  std::condition_variable tick_cv_;
  bool ticking_ = false;
};

}  // namespace plugin
