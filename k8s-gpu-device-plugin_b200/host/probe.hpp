// probe.hpp — probe dispatch for the device map, the ListAndWatch health feed, and the
// per-MIG-slice child process.  All device work goes through the C ABI (include/gpuprobe.h).
//
// Reference sites (uppercaveman/k8s-gpu-device-plugin):
//   device/device_map.go:34-45, :48-98   the build loops that visit each GPU / MIG device
//   device/devices.go:74                 Health = Healthy, replaced by the verdict here
//   plugin/manager.go:156-174            loadPlugins: grade before startPlugins
//   plugin/plugin.go:40,53,173-189       `health` channel and ListAndWatch (no writer today)
#pragma once
#include <fcntl.h>
#include <poll.h>
#include <signal.h>
#include <spawn.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cerrno>
#include <cstring>
#include <deque>
#include <functional>
#include <thread>

#include "device.hpp"

extern char** environ;

namespace plugin {

// What GradeDevices needs from the engine.  GpuProbeBackend is the product; tests substitute
// a scripted backend to exercise the host logic without a GPU.
struct ProbeBackend {
  virtual ~ProbeBackend() = default;
  // full GPUs: all at once, one host thread per device inside the library
  virtual int ProbeMany(const std::vector<std::string>& uuids, uint32_t mask,
                        std::vector<gp_device_result>* out, std::string* err) = 0;
  // one MIG slice: must run in its own process (a CUDA process sees one compute instance)
  // total_mem is the slice's memory (device/device.go:175-181): the probe is sized to it
  virtual int ProbeSlice(const std::string& mig_uuid, uint32_t mask, uint64_t total_mem, gp_device_result* out,
                         std::string* err) = 0;
  // compute processes other than ours on the device right now, -1 when that cannot be told (no NVML):
  // a 16 GiB, full-SM probe perturbs tenants, so busy devices never get the full probe uninvited
  virtual int OtherComputeProcs(const std::string& /*uuid*/) { return -1; }
  // one FULL GPU graded from a fresh child process (CUDA_VISIBLE_DEVICES=<uuid>): a fault in the probe,
  // or a context already lost to a sticky error, cannot touch the plugin's own process
  virtual int ProbeIsolated(const std::string& uuid, uint32_t mask, uint64_t total_mem, gp_device_result* out,
                            std::string* err) { return ProbeSlice(uuid, mask, total_mem, out, err); }
  // the GPU-to-GPU fabric over the first ngpus CUDA devices (K4); uuids = CUDA ordinal order
  virtual int ProbeFabric(int /*ngpus*/, gp_fabric_result* /*out*/, std::vector<std::string>* /*uuids*/, std::string* err) {
    if (err) *err = "fabric probe not available";
    return GP_ERR_UNSUPPORTED;
  }
  // give a kept arena back (Allocate: the pod is about to start) / recover a lost context in-process
  virtual void TrimDevice(const std::string& /*uuid*/) {}
  virtual int ResetDevice(const std::string& /*uuid*/) { return GP_ERR_UNSUPPORTED; }
};

class GpuProbeBackend : public ProbeBackend {
 public:
  explicit GpuProbeBackend(const gp_config* cfg, std::string child_binary = "")
      : child_(std::move(child_binary)) {
    if (cfg) cfg_ = *cfg; else { memset(&cfg_, 0, sizeof(cfg_)); cfg_.struct_size = sizeof(cfg_); cfg_.abi_version = GP_ABI_VERSION; }
    rc_ = gp_init(&cfg_, &ctx_);
    if (rc_) init_error_ = gp_last_error_detail();
  }
  ~GpuProbeBackend() override { if (ctx_) gp_shutdown(ctx_); }
  int init_rc() const { return rc_; }
  const std::string& init_error() const { return init_error_; }
  gp_ctx* ctx() const { return ctx_; }

  int ProbeMany(const std::vector<std::string>& uuids, uint32_t mask, std::vector<gp_device_result>* out,
                std::string* err) override {
    out->assign(uuids.size(), gp_device_result{});
    if (!ctx_) { if (err) *err = "gp_init failed: " + init_error_; return rc_ ? rc_ : GP_ERR_CUDA; }
    std::vector<int> rcs(uuids.size(), 0);
    std::vector<std::string> errs(uuids.size());
    std::vector<std::thread> th;
    for (size_t i = 0; i < uuids.size(); ++i)
      th.emplace_back([&, i] {
        rcs[i] = gp_probe_device(ctx_, uuids[i].c_str(), mask, &(*out)[i]);
        if (rcs[i]) errs[i] = gp_last_error_detail();  // thread-local: read it on this thread
      });
    for (auto& t : th) t.join();
    for (size_t i = 0; i < uuids.size(); ++i)
      if (rcs[i]) { if (err) *err = uuids[i] + ": " + errs[i]; return rcs[i]; }
    return GP_OK;
  }

  int OtherComputeProcs(const std::string& uuid) override {
    gp_nvml_info i;
    if (!ctx_ || gp_nvml_snapshot(ctx_, uuid.c_str(), &i) != GP_OK || !i.available) return -1;
    return (int)i.other_compute_procs;
  }

  int ProbeFabric(int ngpus, gp_fabric_result* out, std::vector<std::string>* uuids, std::string* err) override {
    if (!ctx_) { if (err) *err = "gp_init failed: " + init_error_; return rc_ ? rc_ : GP_ERR_CUDA; }
    const int rc = gp_probe_fabric(ctx_, ngpus, out);
    if (rc) { if (err) *err = gp_last_error_detail(); return rc; }
    if (uuids) {
      uuids->clear();
      for (int i = 0; i < ngpus; ++i) {
        char u[64] = "";
        gp_device_uuid(ctx_, i, u, sizeof(u));
        uuids->push_back(u);
      }
    }
    return GP_OK;
  }
  void TrimDevice(const std::string& uuid) override { if (ctx_) gp_trim_device(ctx_, uuid.c_str()); }
  int ResetDevice(const std::string& uuid) override { return ctx_ ? gp_device_reset(ctx_, uuid.c_str()) : GP_ERR_CUDA; }

  // arrays of a slice probe: the configured size, but never more than a quarter of the slice
  // (triad needs three arrays), rounded down to 1 MiB
  static uint64_t SliceBytes(uint64_t configured, uint64_t total_mem) {
    uint64_t want = configured ? configured : (16ull << 30);
    if (total_mem) want = std::min<uint64_t>(want, total_mem / 4);
    want &= ~((1ull << 20) - 1);
    return std::max<uint64_t>(want, 1ull << 20);
  }

  // how long a child may take before it is killed and the device reported as not answering
  void SetChildTimeout(std::chrono::milliseconds t) { child_timeout_ = t; }
  // extra argv for the child (tests: --test-hang, --test-cuda-error N)
  void SetChildExtraArgs(std::vector<std::string> a) { child_extra_ = std::move(a); }

  // Spawn the slice binary with CUDA_VISIBLE_DEVICES=<uuid>; it writes the POD result struct to its stdout.
  // The plugin is heavily multithreaded (CUDA, gRPC): everything the child needs — argv, envp — is built
  // BEFORE the spawn and the child goes straight to exec (posix_spawn), so nothing between fork and exec can
  // wait on a lock another thread held.  The pipe is close-on-exec, the read has a deadline, and a child
  // that does not answer in time is killed and reaped: the caller always gets control back.
  int ProbeSlice(const std::string& uuid, uint32_t mask, uint64_t total_mem, gp_device_result* out, std::string* err) override {
    if (child_.empty()) { if (err) *err = "no slice binary configured"; return GP_ERR_UNSUPPORTED; }
    const uint64_t hbm = SliceBytes(cfg_.hbm_bytes, total_mem);
    const uint64_t pat = std::min<uint64_t>(cfg_.pattern_bytes ? cfg_.pattern_bytes : hbm, hbm);
    std::vector<std::string> args = {child_, std::to_string(mask), std::to_string(hbm), std::to_string(pat),
                                     std::to_string(cfg_.wo_passes), std::to_string(cfg_.gemm_mnk),
                                     std::to_string(cfg_.iters), std::to_string(cfg_.warmup)};
    for (const auto& a : child_extra_) args.push_back(a);
    std::vector<char*> argv;
    for (auto& a : args) argv.push_back(const_cast<char*>(a.c_str()));
    argv.push_back(nullptr);
    std::vector<std::string> envs;
    for (char** e = environ; e && *e; ++e)
      if (strncmp(*e, "CUDA_VISIBLE_DEVICES=", 21) != 0) envs.emplace_back(*e);
    envs.push_back("CUDA_VISIBLE_DEVICES=" + uuid);
    std::vector<char*> envp;
    for (auto& e : envs) envp.push_back(const_cast<char*>(e.c_str()));
    envp.push_back(nullptr);

    int fds[2];
    if (pipe2(fds, O_CLOEXEC) != 0) { if (err) *err = "pipe2 failed"; return GP_ERR_INVALID; }
    posix_spawn_file_actions_t fa;
    posix_spawn_file_actions_init(&fa);
    posix_spawn_file_actions_adddup2(&fa, fds[1], STDOUT_FILENO);  // dup2 clears close-on-exec on the copy
    pid_t pid = -1;
    const int sp = posix_spawn(&pid, child_.c_str(), &fa, nullptr, argv.data(), envp.data());
    posix_spawn_file_actions_destroy(&fa);
    close(fds[1]);
    if (sp != 0) { close(fds[0]); if (err) *err = "posix_spawn(" + child_ + ") failed: " + strerror(sp); return GP_ERR_INVALID; }

    const auto deadline = std::chrono::steady_clock::now() + child_timeout_;
    size_t got = 0;
    char* dst = reinterpret_cast<char*>(out);
    bool timed_out = false;
    while (got < sizeof(*out)) {
      const auto left = std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now()).count();
      if (left <= 0) { timed_out = true; break; }
      struct pollfd pfd = {fds[0], POLLIN, 0};
      const int pr = poll(&pfd, 1, (int)std::min<long long>(left, 1000));
      if (pr < 0 && errno != EINTR) break;
      if (pr <= 0) continue;
      const ssize_t n = read(fds[0], dst + got, sizeof(*out) - got);
      if (n <= 0) break;  // EOF: the child died before answering
      got += (size_t)n;
    }
    close(fds[0]);
    int status = 0;
    if (timed_out) kill(pid, SIGKILL);
    waitpid(pid, &status, 0);
    if (timed_out) {
      memset(out, 0, sizeof(*out));
      if (err) *err = "probe process for " + uuid + " did not answer within " + std::to_string(child_timeout_.count()) + " ms and was killed";
      return GP_ERR_DEVICE_LOST;
    }
    if (got != sizeof(*out) || out->struct_size != sizeof(*out)) {
      memset(out, 0, sizeof(*out));
      const std::string how = WIFSIGNALED(status) ? "killed by signal " + std::to_string(WTERMSIG(status))
                                                  : "exit status " + std::to_string(WIFEXITED(status) ? WEXITSTATUS(status) : -1);
      if (err) *err = "probe process for " + uuid + " returned no result (" + how + ")";
      return GP_ERR_DEVICE_LOST;
    }
    return out->error;
  }

 private:
  gp_config cfg_;
  gp_ctx* ctx_ = nullptr;
  int rc_ = 0;
  std::string init_error_, child_;
  std::chrono::milliseconds child_timeout_{120000};
  std::vector<std::string> child_extra_;
};

// The new step between building the map and advertising it (device_map.go:34-45 ->
// manager.go:156-174): full GPUs are graded together, MIG slices one child process each, all
// concurrently, never serially inside the NVML visitor.  A device whose probe errors is
// Unhealthy, never silently Healthy.  Returns the number of Unhealthy devices.
// true iff the device's last probe ran clean and passed
inline bool VerdictHealthy(const Device& d) { return d.Probe && d.Probe->error == 0 && d.Probe->healthy == 1; }

// apply_health = true at start-up (nothing is advertised yet, Health is written directly, as
// BuildDevice does); false for a re-probe while ListAndWatch streams are open: then only
// Device.Probe is updated and the caller routes the transition through
// NvidiaDevicePlugin::Report, which owns Health and its mutex.
inline int GradeDevices(DeviceMap& dm, ProbeBackend& be, uint32_t mask, std::vector<std::string>* log = nullptr,
                        bool apply_health = true) {
  std::vector<std::shared_ptr<Device>> gpus, slices;
  for (auto& d : dm.All()) (d->IsMigDevice() ? slices : gpus).push_back(d);
  std::vector<gp_device_result> res;
  std::vector<std::string> uuids;
  for (auto& d : gpus) uuids.push_back(d->GetUUID());
  std::string err;
  int rc = GP_OK;
  std::vector<std::thread> th;
  std::vector<gp_device_result> sres(slices.size());
  std::vector<int> src(slices.size(), 0);
  std::vector<std::string> serr(slices.size());
  for (size_t i = 0; i < slices.size(); ++i)
    th.emplace_back([&, i] { src[i] = be.ProbeSlice(slices[i]->GetUUID(), mask, slices[i]->TotalMemory, &sres[i], &serr[i]); });
  if (!gpus.empty()) rc = be.ProbeMany(uuids, mask, &res, &err);
  for (auto& t : th) t.join();
  int unhealthy = 0;
  auto apply = [&](Device& d, const gp_device_result& r, int prc, const std::string& e) {
    const bool ok = prc == GP_OK && r.error == 0 && r.healthy == 1;
    auto pr = std::make_shared<gp_device_result>(r);
    if (!ok) { pr->healthy = 0; if (pr->error == 0 && prc != GP_OK) pr->error = prc; }
    d.Probe = pr;
    if (apply_health) d.Health = ok ? Healthy : Unhealthy;
    if (!ok) {
      ++unhealthy;
      if (log) log->push_back("'" + d.ID + "' graded unhealthy: rc=" + std::to_string(prc) + " failed_mask=" + std::to_string(r.failed_mask) + (e.empty() ? "" : " (" + e + ")"));
    }
  };
  for (size_t i = 0; i < gpus.size(); ++i) {
    const bool have = i < res.size() && res[i].struct_size == sizeof(gp_device_result);
    apply(*gpus[i], have ? res[i] : gp_device_result{}, have ? res[i].error : (rc ? rc : GP_ERR_CUDA), err);
  }
  for (size_t i = 0; i < slices.size(); ++i) apply(*slices[i], sres[i], src[i], serr[i]);
  // replicas (AnnotatedID uuid::n) share their GPU's verdict
  for (auto& kv : dm.m)
    for (auto& dv : kv.second.m) {
      if (dv.second->Probe) continue;
      for (auto& g : dm.All())
        if (g->GetUUID() == dv.second->GetUUID() && g->Probe) {
          dv.second->Probe = g->Probe;
          if (apply_health) dv.second->Health = g->Health;
        }
    }
  return unhealthy;
}

// plugin/plugin.go:34-42,173-189.  The reference's ListAndWatch sends the list once, then blocks
// on `health`; on receive it sets Unhealthy and resends.  Nothing ever sends.  This mirror keeps
// that loop and adds what is missing: a writer (Report), the way back to Healthy, and a mutex
// around Device.Health (written at plugin.go:182 while devices.go:159-166 reads it).
class NvidiaDevicePlugin {
 public:
  using Sender = std::function<bool(const std::vector<PluginDevice>&)>;  // false = stream closed
  explicit NvidiaDevicePlugin(std::string resourceName, Devices devices)
      : resourceName_(std::move(resourceName)), devices_(std::move(devices)) {}

  const Devices& devices() const { return devices_; }

  // the `health` channel's writer: called by the periodic / on-demand re-probe
  void Report(const std::string& id, bool healthy) {
    std::lock_guard<std::mutex> g(mu_);
    events_.push_back({id, healthy});
    cv_.notify_all();
  }
  void Stop() {
    std::lock_guard<std::mutex> g(mu_);
    stop_ = true;
    cv_.notify_all();
  }
  // blocks until Stop(); returns the number of lists sent
  int ListAndWatch(const Sender& send) {
    int sent = 0;
    if (!send(snapshot())) return sent;
    ++sent;
    for (;;) {
      std::pair<std::string, bool> ev;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return stop_ || !events_.empty(); });
        if (events_.empty()) return sent;  // stop requested and nothing pending
        ev = events_.front();
        events_.pop_front();
        auto d = devices_.GetByID(ev.first);
        if (!d) continue;
        const char* h = ev.second ? Healthy : Unhealthy;
        if (d->Health == h) continue;  // no transition, nothing to tell kubelet
        d->Health = h;
      }
      if (!send(snapshot())) return sent;
      ++sent;
    }
  }
  std::vector<PluginDevice> snapshot() {
    std::lock_guard<std::mutex> g(mu_);
    return devices_.GetPluginDevices();
  }
  // plugin/plugin.go:210-225.  New: the devices kubelet is handing to a pod give their kept probe arena back
  // first (gp_trim_device), so the pod finds the whole memory free.
  void SetOnAllocate(std::function<void(const std::vector<std::string>&)> f) { on_allocate_ = std::move(f); }
  std::pair<std::string, Error> Allocate(const std::vector<std::string>& ids) const {
    if (!devices_.Contains(ids)) return {"", "invalid allocation request for " + resourceName_};
    if (on_allocate_) on_allocate_(ids);
    std::string v;
    for (size_t i = 0; i < ids.size(); ++i) v += (i ? "," : "") + ids[i];
    return {"NVIDIA_VISIBLE_DEVICES=" + v, std::nullopt};
  }

 private:
  std::string resourceName_;
  Devices devices_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<std::pair<std::string, bool>> events_;
  std::function<void(const std::vector<std::string>&)> on_allocate_;
  bool stop_ = false;
};

}  // namespace plugin
