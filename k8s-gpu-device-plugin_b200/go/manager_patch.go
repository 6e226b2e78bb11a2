// NOT COMPILED in this repository (no Go toolchain in the build image).  The patch to
// plugin/manager.go and plugin/plugin.go; host/manager.hpp (PluginManager) is the compiled,
// tested C++ twin, function for function (INTEGRATION.md §5).
package plugin

import (
	"fmt"
	"sort"
	"time"

	bmk "github.com/uppercaveman/k8s-gpu-device-plugin/benchmark"
	"github.com/uppercaveman/k8s-gpu-device-plugin/device"
	"github.com/uppercaveman/k8s-gpu-device-plugin/metrics"
)

// New fields on PluginManager (plugin/manager.go:20-34):
//	probe       *bmk.Probe
//	probeMask   uint32
//	last        map[string]*bmk.ProbeResult   // by uuid, survives restartPlugins
//	fabric      *bmk.FabricResult             // measured P2P matrix for GetPreferredAllocation
//	isolate     bool                          // probe.isolate: periodic re-probes from child processes

// gradeAtLoad is called in loadPlugins (plugin/manager.go:156-174) between NewDeviceMap (:158) and
// the plugin loop (:165).  restartPlugins runs loadPlugins again on every kubelet socket
// re-creation and GET /restart, i.e. on nodes whose GPUs run pods: a device somebody else is
// computing on never gets the full probe here and keeps the verdict of its last grade.  When
// busy-ness cannot be told (no NVML) the probe is tried and GP_ERR_NOMEM means "in use, not
// graded".  Twin: PluginManager::loadPlugins.
func (p *PluginManager) gradeAtLoad(dm device.DeviceMap) {
	if p.probe == nil {
		return
	}
	idle, busy := device.DeviceMap{}, 0
	procs := map[string]int{}
	for name, devs := range dm {
		for id, d := range devs {
			u := d.GetUUID()
			if _, ok := procs[u]; !ok {
				procs[u] = p.probe.OtherComputeProcs(u)
			}
			if procs[u] > 0 {
				busy++
				p.keepLastVerdict(d)
				continue
			}
			if idle[name] == nil {
				idle[name] = device.Devices{}
			}
			idle[name][id] = d
		}
	}
	device.GradeDevices(idle, p.probe, p.probeMask, true)
	for _, devs := range idle {
		for _, d := range devs {
			if procs[d.GetUUID()] < 0 && d.Probe != nil && d.Probe.Error == bmk.ErrNoMem {
				p.keepLastVerdict(d) // a tenant holds the memory: not graded, not unhealthy
			} else if d.Probe != nil {
				p.last[d.GetUUID()] = d.Probe
			}
			if d.Probe != nil {
				metrics.Update(d.Index, *d.Probe)
			}
		}
	}
	if busy == 0 {
		p.gradeFabric(dm)
	}
}

func (p *PluginManager) keepLastVerdict(d *device.Device) {
	if r, ok := p.last[d.GetUUID()]; ok {
		d.Probe = r
		if r.Healthy {
			d.Health = "Healthy"
		} else {
			d.Health = "Unhealthy"
		}
		return
	}
	d.Probe, d.Health = nil, "Healthy"
}

// gradeFabric: K4 once per load, every full GPU idle.  Twin: PluginManager::GradeFabric.
func (p *PluginManager) gradeFabric(dm device.DeviceMap) {
	full := map[string]bool{}
	for _, devs := range dm {
		for _, d := range devs {
			if !d.IsMigDevice() && !device.AnnotatedID(d.ID).HasAnnotations() {
				full[d.GetUUID()] = true
			}
		}
	}
	if len(full) < 2 {
		return
	}
	f, err := p.probe.ProbeFabric(len(full))
	if err != nil {
		return
	}
	p.fabric = &f
	metrics.UpdateFabric(f)
}

// preferredFromMatrix replaces the NVML link query of alignedAlloc (plugin/plugin.go:256-282, whose
// nvmllib is never set): greedily grow `required` with the candidate whose WORST measured link
// to the chosen set is best (ties: lowest id).  Twin: alignedAllocFromMatrix.
func (p *PluginManager) preferredFromMatrix(available, required []string, size int) ([]string, error) {
	f := p.fabric
	ord := func(id string) int {
		if f != nil {
			for i, u := range f.UUIDs {
				if u == id {
					return i
				}
			}
		}
		return -1
	}
	link := func(a, b string) float64 {
		if f == nil {
			return 1 // no matrix yet: flat, i.e. id order
		}
		i, j := ord(a), ord(b)
		if i < 0 || j < 0 {
			return 0
		}
		m := func(x, y int) float64 {
			r, w := f.Read[x*f.N+y], f.Write[x*f.N+y]
			if w < r {
				return w
			}
			return r
		}
		if m(j, i) < m(i, j) {
			return m(j, i)
		}
		return m(i, j)
	}
	chosen := append([]string{}, required...)
	var cand []string
	for _, a := range available {
		in := false
		for _, r := range required {
			in = in || r == a
		}
		if !in {
			cand = append(cand, a)
		}
	}
	sort.Strings(cand)
	if len(chosen)+len(cand) < size {
		return nil, fmt.Errorf("not enough available devices to satisfy allocation")
	}
	for len(chosen) < size {
		best, bestScore := -1, -1.0
		for c := range cand {
			worst := 1e30
			against := chosen
			if len(chosen) == 0 {
				against = cand
			}
			for _, s := range against {
				if s != cand[c] {
					if l := link(cand[c], s); l < worst {
						worst = l
					}
				}
			}
			if worst > bestScore {
				bestScore, best = worst, c
			}
		}
		chosen = append(chosen, cand[best])
		cand = append(cand[:best], cand[best+1:]...)
	}
	return chosen, nil
}

// In Allocate (plugin/plugin.go:210-225), after the Contains check: hand the kept probe arena of
// every requested device back before the pod starts.
//	for _, id := range req.DevicesIDs { plugin.probe.TrimDevice(device.AnnotatedID(id).GetID()) }

// Reprobe re-grades all devices (or one uuid) while ListAndWatch streams are open.  Health is not
// written here: each transition goes to the owning plugin's `health` channel through
// device.Report, which never blocks; the ListAndWatch arm owns Device.Health.  skipBusy is the
// periodic path: devices with tenants keep their last verdict.  A device whose in-process context
// was lost (GP_ERR_DEVICE_LOST), or every full GPU when probe.isolate is set, is graded from a
// fresh child process; if the child finds it healthy the in-process context is reset.
// Twin: PluginManager::Reprobe + GradeIsolated.
func (p *PluginManager) Reprobe(uuid string, skipBusy bool) ([]bmk.ProbeResult, error) {
	var out []bmk.ProbeResult
	seen := map[string]bool{}
	known := false
	for _, pl := range p.plugins {
		np, ok := pl.(*NvidiaDevicePlugin)
		if !ok {
			continue
		}
		for _, d := range np.devices {
			u := d.GetUUID()
			if seen[u] || (uuid != "" && uuid != u) {
				continue
			}
			seen[u] = true
			known = true
			if skipBusy && p.probe.OtherComputeProcs(u) > 0 {
				continue
			}
			lost := d.Probe != nil && d.Probe.Error == bmk.ErrDeviceLost
			var res bmk.ProbeResult
			var err error
			switch {
			case d.IsMigDevice() || p.isolate || lost:
				res, err = p.probe.ProbeChild(u, p.probeMask, d.TotalMemory)
				if err == nil && res.Healthy && lost {
					p.probe.ResetDevice(u)
				}
			default:
				res, err = p.probe.ProbeDevice(u, p.probeMask)
			}
			if err != nil || res.Error != 0 {
				res.Healthy = false
			}
			r := res
			d.Probe = &r
			p.last[u] = &r
			metrics.Update(d.Index, res)
			device.Report(np.health, device.HealthEvent{Device: d, Healthy: res.Healthy})
			out = append(out, res)
		}
	}
	if !known {
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
				_, _ = p.Reprobe("", true)
			}
		}
	}()
}
