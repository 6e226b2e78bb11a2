// device.hpp — C++ mirror of the reference's device model for the probe path.
//
// The reference is Go and there is no Go toolchain in this image, so the host side above the
// C ABI is written in C++ with the reference's names, argument meaning and error text:
//   Device / Devices / AnnotatedID  <- device/devices.go:21-265
//   deviceInfo                      <- device/devices.go:12-19
//   BuildDevice                     <- device/devices.go:41-85   (Health set at :74)
//   DeviceMap / setEntry / wildCardToRegexp <- device/device_map.go:21-125
// What is new is GradeDevices (probe.hpp): the verdict that replaces the hard-coded Healthy.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <optional>
#include <regex>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "../../include/gpuprobe.h"

namespace plugin {

// pluginapi.Healthy / pluginapi.Unhealthy (k8s.io/kubelet deviceplugin/v1beta1 constants)
inline const char* const Healthy = "Healthy";
inline const char* const Unhealthy = "Unhealthy";

// Go returns (value, error); an empty optional<string> error means nil.
using Error = std::optional<std::string>;

// device/devices.go:12-19
struct deviceInfo {
  virtual ~deviceInfo() = default;
  virtual std::pair<std::string, Error> GetUUID() = 0;
  virtual std::pair<std::vector<std::string>, Error> GetPaths() = 0;
  // (hasNuma, node, err)
  virtual std::tuple<bool, int, Error> GetNumaNode() = 0;
  virtual std::pair<uint64_t, Error> GetTotalMemory() = 0;
  virtual std::pair<std::string, Error> GetComputeCapability() = 0;
};

// pluginapi.Device: the part kubelet sees
struct PluginDevice {
  std::string ID;
  std::string Health;
  std::optional<int64_t> NumaNode;  // Topology.Nodes[0].ID
};

// device/devices.go:21-29, plus the probe result (never serialised to kubelet)
struct Device : PluginDevice {
  std::vector<std::string> Paths;
  std::string Index;
  uint64_t TotalMemory = 0;
  std::string ComputeCapability;
  int Replicas = 0;
  std::shared_ptr<gp_device_result> Probe;  // last verdict, nullptr until graded

  bool IsMigDevice() const { return Index.find(':') != std::string::npos; }  // devices.go:212
  bool AlignedAllocationSupported() const {                                   // devices.go:197-209
    if (IsMigDevice()) return false;
    for (const auto& p : Paths)
      if (p == "/dev/dxg") return false;
    return true;
  }
  std::string GetUUID() const;  // devices.go:217
};

// device/devices.go:222-265
struct AnnotatedID {
  std::string s;
  static AnnotatedID New(const std::string& id, int replica) { return {id + "::" + std::to_string(replica)}; }
  bool HasAnnotations() const { return s.find("::") != std::string::npos; }
  std::pair<std::string, int> Split() const {
    const auto p = s.find("::");
    if (p == std::string::npos) return {s, 0};
    int replica = 0;
    try { replica = std::stoi(s.substr(p + 2)); } catch (...) { replica = 0; }
    return {s.substr(0, p), replica};
  }
  std::string GetID() const { return Split().first; }
};

inline std::string Device::GetUUID() const { return AnnotatedID{ID}.GetID(); }

// device/devices.go:32 — map[string]*Device with the set algebra the plugin uses
class Devices {
 public:
  using Map = std::map<std::string, std::shared_ptr<Device>>;
  Map m;

  bool Contains(const std::vector<std::string>& ids) const {  // :88-95
    for (const auto& id : ids)
      if (!m.count(id)) return false;
    return true;
  }
  std::shared_ptr<Device> GetByID(const std::string& id) const {  // :98
    auto it = m.find(id);
    return it == m.end() ? nullptr : it->second;
  }
  std::shared_ptr<Device> GetByIndex(const std::string& index) const {  // :103-110
    for (const auto& kv : m)
      if (kv.second->Index == index) return kv.second;
    return nullptr;
  }
  Devices Subset(const std::vector<std::string>& ids) const {  // :113-121
    Devices r;
    for (const auto& id : ids)
      if (m.count(id)) r.m[id] = m.at(id);
    return r;
  }
  Devices Difference(const Devices& o) const {  // :124-132
    Devices r;
    for (const auto& kv : m)
      if (!o.m.count(kv.first)) r.m[kv.first] = kv.second;
    return r;
  }
  std::vector<std::string> GetIDs() const {  // :135
    std::vector<std::string> r;
    for (const auto& kv : m) r.push_back(kv.second->ID);
    return r;
  }
  std::vector<std::string> GetUUIDs() const {  // :144-156 (replicas collapse to one uuid)
    std::vector<std::string> r;
    std::set<std::string> seen;
    for (const auto& kv : m) {
      const std::string u = kv.second->GetUUID();
      if (seen.insert(u).second) r.push_back(u);
    }
    return r;
  }
  // :159-166 — COPIES each device, so Health must be set on the map entry before this is called
  std::vector<PluginDevice> GetPluginDevices() const {
    std::vector<PluginDevice> r;
    for (const auto& kv : m) r.push_back(static_cast<const PluginDevice&>(*kv.second));
    return r;
  }
  std::vector<std::string> GetIndices() const {  // :169
    std::vector<std::string> r;
    for (const auto& kv : m) r.push_back(kv.second->Index);
    return r;
  }
  bool AlignedAllocationSupported() const {  // :187-194
    for (const auto& kv : m)
      if (!kv.second->AlignedAllocationSupported()) return false;
    return true;
  }
  size_t size() const { return m.size(); }
};

// device/devices.go:41-85.  Same five getters in the same order, same error prefixes; the one
// behavioural difference is deliberate: Health starts as Healthy exactly as at :74 and is then
// overwritten by GradeDevices before anything is advertised.
inline std::pair<std::shared_ptr<Device>, Error> BuildDevice(const std::string& index, deviceInfo& d) {
  auto [uuid, e1] = d.GetUUID();
  if (e1) return {nullptr, "error getting UUID device: " + *e1};
  auto [paths, e2] = d.GetPaths();
  if (e2) return {nullptr, "error getting device paths: " + *e2};
  auto [hasNuma, numa, e3] = d.GetNumaNode();
  if (e3) return {nullptr, "error getting device NUMA node: " + *e3};
  auto [totalMemory, e4] = d.GetTotalMemory();
  if (e4) return {nullptr, "error getting device memory: " + *e4};
  auto [cc, e5] = d.GetComputeCapability();
  if (e5) return {nullptr, "error getting device compute capability: " + *e5};
  auto dev = std::make_shared<Device>();
  dev->TotalMemory = totalMemory;
  dev->ComputeCapability = cc;
  dev->ID = uuid;
  dev->Index = index;
  dev->Paths = paths;
  dev->Health = Healthy;
  if (hasNuma) dev->NumaNode = numa;
  return {dev, std::nullopt};
}

// device/device_map.go:114-125
inline std::string wildCardToRegexp(const std::string& pattern) {
  std::string out;
  size_t start = 0;
  bool first = true;
  static const std::string meta = R"(\.+*?()|[]{}^$)";
  for (;;) {
    const size_t star = pattern.find('*', start);
    const std::string lit = pattern.substr(start, star == std::string::npos ? std::string::npos : star - start);
    if (!first) out += ".*";
    for (char c : lit) {
      if (meta.find(c) != std::string::npos) out += '\\';
      out += c;
    }
    first = false;
    if (star == std::string::npos) break;
    start = star + 1;
  }
  return out;
}

// device/device_map.go:21 — resource name -> Devices
class DeviceMap {
 public:
  std::map<std::string, Devices> m;
  // :101-111
  Error setEntry(const std::string& name, const std::string& index, deviceInfo& info) {
    auto [dev, err] = BuildDevice(index, info);
    if (err) return "error building Device: " + *err;
    m[name].m[dev->ID] = dev;
    return std::nullopt;
  }
  // every device across resources, replicas collapsed (what the probe iterates over)
  std::vector<std::shared_ptr<Device>> All() const {
    std::vector<std::shared_ptr<Device>> r;
    std::set<std::string> seen;
    for (const auto& kv : m)
      for (const auto& d : kv.second.m)
        if (seen.insert(d.second->GetUUID() + (d.second->IsMigDevice() ? "|mig" : "|gpu")).second) r.push_back(d.second);
    return r;
  }
};

// resource pattern matching of device_map.go:59-69: Go's regexp.MatchString is unanchored
inline bool matchResourcePattern(const std::string& pattern, const std::string& name) {
  return std::regex_search(name, std::regex(wildCardToRegexp(pattern)));
}

}  // namespace plugin
