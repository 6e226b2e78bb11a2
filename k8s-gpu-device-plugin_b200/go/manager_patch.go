// NOT COMPILED in this repository (no Go toolchain in the build image).  The patch to
// plugin/manager.go; host/manager.hpp is the compiled, tested C++ twin.
package plugin

import (
	"fmt"
	"time"

	bmk "github.com/uppercaveman/k8s-gpu-device-plugin/benchmark"
	"github.com/uppercaveman/k8s-gpu-device-plugin/device"
	"github.com/uppercaveman/k8s-gpu-device-plugin/metrics"
)

// New fields on PluginManager (plugin/manager.go:20-34):
//	probe     *bmk.Probe
//	probeMask uint32

// In loadPlugins (plugin/manager.go:156-174), between NewDeviceMap (:158) and the plugin loop (:165):
//
//	if p.probe != nil {
//	    if n := device.GradeDevices(dmp, p.probe, p.probeMask); n > 0 {
//	        l.Logger.Warn("devices graded unhealthy at start-up", zap.Int("count", n))
//	    }
//	    p.probe.Trim() // pods need the memory
//	}

// Reprobe re-grades all devices (or one uuid) while ListAndWatch streams are open.  Health is not
// written here: each transition is sent on the owning plugin's `health` channel, whose
// ListAndWatch arm (plugin/plugin.go:181-186, patched in go/grade.go) owns Device.Health.
func (p *PluginManager) Reprobe(uuid string) ([]bmk.ProbeResult, error) {
	var out []bmk.ProbeResult
	seen := map[string]bool{}
	for _, pl := range p.plugins {
		np, ok := pl.(*NvidiaDevicePlugin)
		if !ok {
			continue
		}
		for _, d := range np.devices {
			u := d.GetUUID()
			if d.IsMigDevice() || seen[u] || (uuid != "" && uuid != u) {
				continue
			}
			seen[u] = true
			res, err := p.probe.ProbeDevice(u, p.probeMask)
			metrics.Update(d.Index, res)
			np.health <- device.HealthEvent{Device: d, Healthy: err == nil && res.Healthy}
			out = append(out, res)
		}
	}
	p.probe.Trim()
	if len(out) == 0 {
		return nil, fmt.Errorf("unknown device %s", uuid)
	}
	return out, nil
}

// StartProbeTicker is the periodic feed of plugin.health (SURVEY.md §8f.1).
func (p *PluginManager) StartProbeTicker(every time.Duration) {
	go func() {
		t := time.NewTicker(every)
		defer t.Stop()
		for {
			select {
			case <-p.ctx.Done():
				return
			case <-t.C:
				_, _ = p.Reprobe("")
			}
		}
	}()
}
