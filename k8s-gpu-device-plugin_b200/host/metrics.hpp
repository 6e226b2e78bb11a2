// metrics.hpp — the gauges the empty reference package (metrics/metrics.go:1, `package metrics`)
// gains.  The Go side registers them on the Prometheus default registry so they appear on the
// existing GET /metrics (router/api.go:32, promhttp.Handler()); this mirror renders the same
// families in the text exposition format, which is what that handler emits.
#pragma once
#include <cstdio>
#include <map>
#include <mutex>
#include <sstream>
#include <string>
#include <vector>

#include "device.hpp"

namespace plugin {

class ProbeMetrics {
 public:
  struct Family { std::string help; std::map<std::string, double> series; };  // label string -> value

  void Set(const std::string& name, const std::string& help, const std::string& labels, double v) {
    std::lock_guard<std::mutex> g(mu_);
    auto& f = fam_[name];
    f.help = help;
    f.series[labels] = v;
  }
  // called after every probe (startup grade, periodic tick, GET /probe)
  void Update(const Device& d) {
    if (!d.Probe) return;
    const gp_device_result& r = *d.Probe;
    const std::string l = "index=\"" + d.Index + "\",uuid=\"" + d.GetUUID() + "\"";
    Set("gpu_probe_healthy", "1 if every enabled probe passed on the last run", l, r.healthy);
    Set("gpu_probe_duration_seconds", "Wall time of the last gp_probe_device call", l, r.duration_s);
    if (r.probed_mask & GP_PROBE_COPY) Set("gpu_probe_hbm_copy_gbps", "HBM copy bandwidth, algorithmic bytes 2N, median", l, r.copy_gbps_median);
    if (r.probed_mask & GP_PROBE_TRIAD) Set("gpu_probe_hbm_triad_gbps", "HBM triad bandwidth, algorithmic bytes 3N, median", l, r.triad_gbps_median);
    if (r.probed_mask & GP_PROBE_PATTERN) {
      Set("gpu_probe_hbm_fill_gbps", "Pattern fill bandwidth, algorithmic bytes N, median", l, r.fill_gbps_median);
      Set("gpu_probe_hbm_verify_gbps", "Pattern verify bandwidth, algorithmic bytes N, median", l, r.verify_gbps_median);
      Set("gpu_probe_pattern_mismatches", "Words that differed from the expected pattern on the last run", l, (double)r.pattern_mismatches);
      Set("gpu_probe_pattern_passes", "Fill+verify passes run", l, r.pattern_passes);
    }
    if (r.probed_mask & GP_PROBE_GEMM) {
      Set("gpu_probe_gemm_tflops", "bf16 GEMM throughput on tcgen05, median", l, r.gemm.tflops_median);
      Set("gpu_probe_gemm_max_scaled_err", "max |c-ref| / sum|a||b| over the sampled entries", l, r.gemm.max_scaled_err);
    }
  }
  void SetP2P(int src, int dst, const char* dir, double gbps) {
    Set("gpu_probe_p2p_gbps", "NVLink peer bandwidth measured with in-kernel loads/stores",
        "dir=\"" + std::string(dir) + "\",dst=\"" + std::to_string(dst) + "\",src=\"" + std::to_string(src) + "\"", gbps);
  }
  void SetAllReduce(int nranks, double busbw) {
    Set("gpu_probe_allreduce_busbw_gbps", "ncclAllReduce bus bandwidth", "ranks=\"" + std::to_string(nranks) + "\"", busbw);
  }
  // Prometheus text exposition format 0.0.4
  std::string Render() {
    std::lock_guard<std::mutex> g(mu_);
    std::ostringstream o;
    for (const auto& f : fam_) {
      o << "# HELP " << f.first << " " << f.second.help << "\n# TYPE " << f.first << " gauge\n";
      for (const auto& s : f.second.series) {
        char buf[64];
        snprintf(buf, sizeof(buf), "%.17g", s.second);
        o << f.first << "{" << s.first << "} " << buf << "\n";
      }
    }
    return o.str();
  }

 private:
  std::mutex mu_;
  std::map<std::string, Family> fam_;
};

}  // namespace plugin
