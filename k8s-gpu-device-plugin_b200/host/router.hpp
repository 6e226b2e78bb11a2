// router.hpp — the HTTP handlers of router/api.go as pure functions (the echo server,
// CORS and middleware of server/ and middleware/ are transport and out of scope).
//
//   GET /         -> Version   router/api.go:40-42
//   GET /metrics  -> promhttp  router/api.go:32        (+ the probe gauges)
//   GET /health   -> Health    router/api.go:45-47
//   GET /restart  -> Restart   router/api.go:50-54
//   GET /probe    -> NEW: re-probe on demand, same shape as Restart, optional ?uuid=
// The JSON envelope {code,data,msg} is modules/util/http.go:3-15; field order as Go marshals it.
#pragma once
#include <functional>
#include <sstream>
#include <string>

#include "metrics.hpp"
#include "probe.hpp"

namespace plugin {

inline std::string jsonEscape(const std::string& s) {
  std::string o;
  for (unsigned char c : s) {
    switch (c) {
      case '"': o += "\\\""; break;
      case '\\': o += "\\\\"; break;
      case '\n': o += "\\n"; break;
      case '\r': o += "\\r"; break;
      case '\t': o += "\\t"; break;
      default:
        if (c < 0x20) { char b[8]; snprintf(b, sizeof(b), "\\u%04x", c); o += b; }
        else o += (char)c;
    }
  }
  return o;
}
// util.Success(data) / util.Failed(code, msg); `data` is already-encoded JSON
inline std::string Success(const std::string& dataJson) { return "{\"code\":0,\"data\":" + dataJson + ",\"msg\":\"success\"}"; }
inline std::string Failed(int code, const std::string& msg) { return "{\"code\":" + std::to_string(code) + ",\"data\":null,\"msg\":\"" + jsonEscape(msg) + "\"}"; }

struct Response { int status; std::string body; };

inline std::string resultJson(const Device& d) {
  std::ostringstream o;
  o.precision(17);
  o << "{\"uuid\":\"" << jsonEscape(d.GetUUID()) << "\",\"index\":\"" << jsonEscape(d.Index) << "\",\"health\":\"" << d.Health << "\"";
  if (d.Probe) {
    const gp_device_result& r = *d.Probe;
    o << ",\"healthy\":" << (r.healthy ? "true" : "false") << ",\"error\":" << r.error << ",\"failed_mask\":" << r.failed_mask
      << ",\"probed_mask\":" << r.probed_mask << ",\"duration_s\":" << r.duration_s << ",\"copy_gbps\":" << r.copy_gbps_median
      << ",\"triad_gbps\":" << r.triad_gbps_median << ",\"fill_gbps\":" << r.fill_gbps_median << ",\"verify_gbps\":" << r.verify_gbps_median
      << ",\"pattern_passes\":" << r.pattern_passes << ",\"pattern_mismatches\":" << r.pattern_mismatches
      << ",\"pattern_digest\":\"" << std::hex << r.pattern_digest << std::dec << "\""
      << ",\"gemm_tflops\":" << r.gemm.tflops_median << ",\"gemm_max_scaled_err\":" << r.gemm.max_scaled_err;
  }
  o << "}";
  return o.str();
}

class API {
 public:
  API(DeviceMap* dm, ProbeBackend* be, ProbeMetrics* metrics, uint32_t mask,
      std::function<void(const std::string& id, bool healthy)> report, std::function<void()> restart)
      : dm_(dm), be_(be), metrics_(metrics), mask_(mask), report_(std::move(report)), restart_(std::move(restart)) {}

  Response Version() const { return {200, Success("\"version : 0.0.1\"")}; }  // modules/version/version.go:4
  Response Health() const { return {200, Success("\"ok\"")}; }
  Response Restart() const { if (restart_) restart_(); return {200, Success("\"ok\"")}; }
  Response Metrics() const { return {200, metrics_->Render()}; }

  // GET /probe[?uuid=GPU-...]: re-grade now, feed ListAndWatch and the gauges, answer with results
  Response Probe(const std::string& uuid = "") {
    std::lock_guard<std::mutex> g(mu_);  // one re-probe at a time, like one restart at a time
    DeviceMap scope;
    bool found = uuid.empty();
    for (auto& kv : dm_->m)
      for (auto& dv : kv.second.m)
        if (uuid.empty() || dv.second->GetUUID() == uuid) { scope.m[kv.first].m[dv.first] = dv.second; found = true; }
    if (!found) return {200, Failed(404, "unknown device " + uuid)};
    GradeDevices(scope, *be_, mask_, nullptr, /*apply_health=*/report_ == nullptr);
    std::string arr = "[";
    bool first = true;
    for (auto& d : scope.All()) {
      metrics_->Update(*d);
      if (report_) report_(d->ID, VerdictHealthy(*d));  // ListAndWatch owns the Health transition
      arr += (first ? "" : ",") + resultJson(*d);
      first = false;
    }
    arr += "]";
    return {200, Success(arr)};
  }

 private:
  DeviceMap* dm_;
  ProbeBackend* be_;
  ProbeMetrics* metrics_;
  uint32_t mask_;
  std::function<void(const std::string&, bool)> report_;
  std::function<void()> restart_;
  std::mutex mu_;
};

}  // namespace plugin
