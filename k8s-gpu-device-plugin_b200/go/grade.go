// NOT COMPILED in this repository (no Go toolchain in the build image).  The patch to the
// reference's device and plugin packages; each block names the file:line it replaces or
// extends.  host/probe.hpp (GradeDevices, NvidiaDevicePlugin) is the compiled, tested C++ twin.
package device

import (
	"sync"

	bmk "github.com/uppercaveman/k8s-gpu-device-plugin/benchmark"
	pluginapi "k8s.io/kubelet/pkg/apis/deviceplugin/v1beta1"
)

// healthMu guards Device.Health: today plugin/plugin.go:182 writes it from the ListAndWatch
// goroutine while device/devices.go:159-166 (GetPluginDevices) reads it.
var healthMu sync.RWMutex

// New field on Device (device/devices.go:21-29), never serialised to kubelet:
//	Probe *bmk.ProbeResult

// GradeDevices runs after build() (device/device_map.go:34-45) and before startPlugins
// (plugin/manager.go:71): the verdict replaces the Healthy that BuildDevice hard-codes at
// device/devices.go:74.  Full GPUs are graded in this process, one goroutine each; every MIG
// slice ("i:j" index, device/device.go:32-34) is graded by one gpuprobe-slice child process,
// sized to the slice's TotalMemory.  A device whose probe errors — a slice with no slice binary
// configured included — is Unhealthy, never silently Healthy.  applyHealth is true at start-up
// (nothing is advertised yet); false for a re-probe, where ListAndWatch owns the transition.
func GradeDevices(dm DeviceMap, p *bmk.Probe, mask uint32, applyHealth bool) (unhealthy int) {
	type job struct {
		d   *Device
		mig bool
	}
	seen := map[string]bool{}
	var jobs []job
	for _, devs := range dm {
		for _, d := range devs {
			if seen[d.GetUUID()] {
				continue
			}
			seen[d.GetUUID()] = true
			jobs = append(jobs, job{d, d.IsMigDevice()})
		}
	}
	verdict := map[string]*bmk.ProbeResult{}
	var wg sync.WaitGroup
	var mu sync.Mutex
	for _, j := range jobs {
		wg.Add(1)
		go func(j job) {
			defer wg.Done()
			var res bmk.ProbeResult
			var err error
			if j.mig {
				res, err = p.ProbeChild(j.d.GetUUID(), mask, j.d.TotalMemory)
			} else {
				res, err = p.ProbeDevice(j.d.GetUUID(), mask)
			}
			if err != nil || res.Error != 0 {
				res.Healthy = false
			}
			mu.Lock()
			verdict[j.d.GetUUID()] = &res
			mu.Unlock()
		}(j)
	}
	wg.Wait()
	healthMu.Lock()
	defer healthMu.Unlock()
	for _, devs := range dm {
		for _, d := range devs { // replicas (uuid::n) share their GPU's verdict
			res, graded := verdict[d.GetUUID()]
			if !graded {
				continue
			}
			d.Probe = res
			if !res.Healthy {
				unhealthy++
			}
			if applyHealth {
				if res.Healthy {
					d.Health = pluginapi.Healthy
				} else {
					d.Health = pluginapi.Unhealthy
				}
			}
		}
	}
	return unhealthy
}

// HealthEvent is what the probe ticker / GET /probe delivers to ListAndWatch.  The reference's
// channel is `chan *device.Device`, unbuffered, and can only mark Unhealthy
// (plugin/plugin.go:40,53,181-186).  The patch:
//
//	plugin/plugin.go:40   health chan device.HealthEvent
//	plugin/plugin.go:53   health: make(chan device.HealthEvent, 4*len(devices)+16),   // buffered
//	plugin/plugin.go:181  case ev := <-plugin.health:
//	                          healthMu.Lock()
//	                          changed := ev.Device.Health != ev.Health()
//	                          ev.Device.Health = ev.Health()
//	                          healthMu.Unlock()
//	                          if changed {
//	                              s.Send(&pluginapi.ListAndWatchResponse{Devices: plugin.Devices().GetPluginDevices()})
//	                          }
//
// and the writer never blocks (Report below): with no ListAndWatch stream open — kubelet not yet
// connected, or between reconnects — an unbuffered send would park the probe ticker and
// GET /probe for ever.  The C++ twin uses an unbounded queue (NvidiaDevicePlugin::Report).
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

// ApplyHealth is the ListAndWatch arm's half (go/startup_fixes.patch, plugin/plugin.go:181): write the verdict
// under the mutex, report whether it was a transition.  Twin: the locked block of NvidiaDevicePlugin::ListAndWatch.
func ApplyHealth(ev HealthEvent) (changed bool) {
	healthMu.Lock()
	defer healthMu.Unlock()
	changed = ev.Device.Health != ev.Health()
	ev.Device.Health = ev.Health()
	return changed
}

// Report delivers ev without ever blocking.  If the buffer is full (nobody has been listening for
// a long time) the verdict is applied to the device directly, so the first list of the next
// ListAndWatch stream (plugin/plugin.go:174) already carries it.
func Report(ch chan HealthEvent, ev HealthEvent) {
	select {
	case ch <- ev:
	default:
		healthMu.Lock()
		ev.Device.Health = ev.Health()
		healthMu.Unlock()
	}
}
