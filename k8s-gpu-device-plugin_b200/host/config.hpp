// config.hpp — the reference's configuration surface plus the probe's optional keys.
//   Config / SetDefaultConfig   <- config/config.go:9-22   (defaults: webListenAddress "9002",
//                                  migStrategy "none", benchmark false, log.level "debug")
//   config.yml                  <- config.yml:1-13
// The schema does not change: an unmodified config.yml loads exactly as before.  The probe adds
// OPTIONAL keys under `probe:` with defaults, so a file that never mentions them is still valid.
// (The reference reads YAML through viper; this is a reader for the two-level subset that file
// uses — scalars and one level of nesting — which is all the schema has.)
#pragma once
#include <cstdint>
#include <map>
#include <sstream>
#include <string>

#include "../../include/gpuprobe.h"

namespace plugin {

struct LogConfig { std::string Level = "debug", FileDir; };
struct ProbeConfig {
  bool Enabled = true;
  uint64_t HbmBytes = 16ull << 30;  // BASELINE.json configs[1]
  uint64_t PatternBytes = 0;        // 0 = same as HbmBytes; GP_PATTERN_ALL_FREE ("all" in the file) = whole memory
  uint32_t WoPasses = 64;
  uint32_t GemmMNK = 8192;
  std::string GemmOutput = "fp32";  // "fp32" (strictest check) or "bf16" (what cuBLAS bf16 peaks write)
  double MinCopyGbps = 0, MinTriadGbps = 0, MinGemmTflops = 0;  // 0 = floor not enforced
  int IntervalSeconds = 0;                                       // 0 = no periodic re-probe
  // fabric (K4): measured once per loadPlugins when every full GPU is idle; floors 0 = not enforced
  bool Fabric = true;
  double MinP2PGbps = 0, MaxP2PSpread = 0, MinAllreduceBusbwGbps = 0;
  uint64_t FabricBytes = 0;          // 0 = 1 GiB per transfer
  bool KeepArena = false;            // keep the arena of an idle device mapped between re-probes (freed in Allocate)
  bool Isolate = false;              // every periodic re-probe of a full GPU from a child process
  uint64_t TriadExpectSum = 0, TriadExpectXor = 0;  // known answer of the triad at HbmBytes (hex in the file); 0 0 = none
};
struct Config {
  std::string WebListenAddress = "9002";  // config.go:17 (no colon — the reference's own default)
  std::string MigStrategy = "none";
  bool Benchmark = false;                 // still means "run the pprof wrapper" (main.go:140-154)
  LogConfig Log;
  ProbeConfig Probe;

  gp_config ToGpConfig() const {
    gp_config c{};
    c.struct_size = sizeof(c);
    c.abi_version = GP_ABI_VERSION;
    c.hbm_bytes = Probe.HbmBytes;
    c.pattern_bytes = Probe.PatternBytes;
    c.wo_passes = Probe.WoPasses;
    c.gemm_mnk = Probe.GemmMNK;
    c.gemm_output = Probe.GemmOutput == "bf16" ? 1u : 0u;
    c.min_copy_gbps = Probe.MinCopyGbps;
    c.min_triad_gbps = Probe.MinTriadGbps;
    c.min_gemm_tflops = Probe.MinGemmTflops;
    c.min_p2p_gbps = Probe.MinP2PGbps;
    c.max_p2p_spread = Probe.MaxP2PSpread;
    c.min_allreduce_busbw_gbps = Probe.MinAllreduceBusbwGbps;
    c.fabric_bytes = Probe.FabricBytes;
    c.keep_arena = Probe.KeepArena ? 1u : 0u;
    c.triad_expect_sum = Probe.TriadExpectSum;
    c.triad_expect_xor = Probe.TriadExpectXor;
    return c;
  }
};

inline std::string trimYaml(std::string s) {
  const auto h = s.find('#');
  if (h != std::string::npos) s = s.substr(0, h);
  while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\r')) s.pop_back();
  size_t i = 0;
  while (i < s.size() && (s[i] == ' ' || s[i] == '\t')) ++i;
  s = s.substr(i);
  if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\''))) s = s.substr(1, s.size() - 2);
  return s;
}

// flat "a.b" -> value map of a two-level YAML document
inline std::map<std::string, std::string> parseYamlSubset(const std::string& text) {
  std::map<std::string, std::string> kv;
  std::istringstream in(text);
  std::string line, section;
  while (std::getline(in, line)) {
    const bool indented = !line.empty() && (line[0] == ' ' || line[0] == '\t');
    const std::string t = trimYaml(line);
    if (t.empty()) continue;
    const auto c = t.find(':');
    if (c == std::string::npos) continue;
    const std::string key = trimYaml(t.substr(0, c)), val = trimYaml(t.substr(c + 1));
    if (!indented) {
      if (val.empty()) { section = key; continue; }
      section.clear();
      kv[key] = val;
    } else if (!section.empty()) {
      kv[section + "." + key] = val;
    }
  }
  return kv;
}

inline Config LoadConfig(const std::string& yamlText) {
  Config c;  // SetDefaultConfig
  const auto kv = parseYamlSubset(yamlText);
  auto get = [&](const char* k) -> const std::string* { auto it = kv.find(k); return it == kv.end() ? nullptr : &it->second; };
  auto truthy = [](const std::string& v) { return v == "true" || v == "True" || v == "yes" || v == "1"; };
  if (auto v = get("webListenAddress")) c.WebListenAddress = *v;
  if (auto v = get("migStrategy")) c.MigStrategy = *v;
  if (auto v = get("benchmark")) c.Benchmark = truthy(*v);
  if (auto v = get("log.level")) c.Log.Level = *v;
  if (auto v = get("log.fileDir")) c.Log.FileDir = *v;
  if (auto v = get("probe.enabled")) c.Probe.Enabled = truthy(*v);
  if (auto v = get("probe.hbmBytes")) c.Probe.HbmBytes = std::stoull(*v);
  if (auto v = get("probe.patternBytes")) c.Probe.PatternBytes = (*v == "all") ? GP_PATTERN_ALL_FREE : std::stoull(*v);
  if (auto v = get("probe.woPasses")) c.Probe.WoPasses = (uint32_t)std::stoul(*v);
  if (auto v = get("probe.gemmMNK")) c.Probe.GemmMNK = (uint32_t)std::stoul(*v);
  if (auto v = get("probe.gemmOutput")) c.Probe.GemmOutput = *v;
  if (auto v = get("probe.minCopyGbps")) c.Probe.MinCopyGbps = std::stod(*v);
  if (auto v = get("probe.minTriadGbps")) c.Probe.MinTriadGbps = std::stod(*v);
  if (auto v = get("probe.minGemmTflops")) c.Probe.MinGemmTflops = std::stod(*v);
  if (auto v = get("probe.intervalSeconds")) c.Probe.IntervalSeconds = std::stoi(*v);
  if (auto v = get("probe.fabric")) c.Probe.Fabric = truthy(*v);
  if (auto v = get("probe.minP2PGbps")) c.Probe.MinP2PGbps = std::stod(*v);
  if (auto v = get("probe.maxP2PSpread")) c.Probe.MaxP2PSpread = std::stod(*v);
  if (auto v = get("probe.minAllreduceBusbwGbps")) c.Probe.MinAllreduceBusbwGbps = std::stod(*v);
  if (auto v = get("probe.fabricBytes")) c.Probe.FabricBytes = std::stoull(*v);
  if (auto v = get("probe.keepArena")) c.Probe.KeepArena = truthy(*v);
  if (auto v = get("probe.isolate")) c.Probe.Isolate = truthy(*v);
  if (auto v = get("probe.triadExpectSum")) c.Probe.TriadExpectSum = std::stoull(*v, nullptr, 0);
  if (auto v = get("probe.triadExpectXor")) c.Probe.TriadExpectXor = std::stoull(*v, nullptr, 0);
  return c;
}

}  // namespace plugin
