// NOT COMPILED in this repository (no Go toolchain in the build image).  The patch to the
// reference's device and plugin packages; each block names the file:line it replaces or
// extends.  host/probe.hpp is the compiled, tested C++ twin of this logic.
package device

import (
	"sync"

	bmk "github.com/uppercaveman/k8s-gpu-device-plugin/benchmark"
	pluginapi "k8s.io/kubelet/pkg/apis/deviceplugin/v1beta1"
)

// healthMu guards Device.Health: today plugin/plugin.go:182 writes it from the ListAndWatch
// goroutine while device/devices.go:159-166 (GetPluginDevices) reads it.
var healthMu sync.RWMutex

// GradeDevices runs after build() (device/device_map.go:34-45) and before startPlugins
// (plugin/manager.go:71): the verdict replaces the Healthy that BuildDevice hard-codes at
// device/devices.go:74.  Full GPUs are graded together; MIG slices ("i:j" indices,
// device/device.go:32-34) need one child process each and are handled by the caller via
// `gpuprobe-slice` (host/probe_slice.cpp).
func GradeDevices(dm DeviceMap, p *bmk.Probe, mask uint32) (unhealthy int) {
	seen := map[string]bool{}
	var uuids []string
	for _, devs := range dm {
		for _, d := range devs {
			if d.IsMigDevice() || seen[d.GetUUID()] {
				continue
			}
			seen[d.GetUUID()] = true
			uuids = append(uuids, d.GetUUID())
		}
	}
	verdict := map[string]bool{}
	var wg sync.WaitGroup
	var mu sync.Mutex
	for _, u := range uuids {
		wg.Add(1)
		go func(u string) {
			defer wg.Done()
			res, err := p.ProbeDevice(u, mask)
			mu.Lock()
			verdict[u] = err == nil && res.Healthy
			mu.Unlock()
		}(u)
	}
	wg.Wait()
	healthMu.Lock()
	defer healthMu.Unlock()
	for _, devs := range dm {
		for _, d := range devs {
			ok, graded := verdict[d.GetUUID()]
			if !graded {
				continue
			}
			if ok {
				d.Health = pluginapi.Healthy
			} else {
				d.Health = pluginapi.Unhealthy // a probe that errored is never silently Healthy
				unhealthy++
			}
		}
	}
	return unhealthy
}

// HealthEvent is what the probe ticker / GET /probe sends on plugin.health.  The reference's
// channel carries *Device and can only mark Unhealthy (plugin/plugin.go:181-186); the patched
// ListAndWatch arm is:
//
//	case ev := <-plugin.health:
//	    healthMu.Lock()
//	    changed := ev.Device.Health != ev.Health()
//	    ev.Device.Health = ev.Health()
//	    healthMu.Unlock()
//	    if changed { s.Send(&pluginapi.ListAndWatchResponse{Devices: plugin.Devices().GetPluginDevices()}) }
type HealthEvent struct {
	Device  *Device
	Healthy bool
}

func (e HealthEvent) Health() string {
	if e.Healthy {
		return pluginapi.Healthy
	}
	return pluginapi.Unhealthy
}
