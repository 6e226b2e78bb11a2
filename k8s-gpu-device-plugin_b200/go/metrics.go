// NOT COMPILED in this repository (no Go toolchain in the build image).  Replaces the empty
// metrics/metrics.go:1 (`package metrics`).  Gauges register on the default registry, so they
// show up on the existing GET /metrics (router/api.go:32) with no route change.
// host/metrics.hpp renders the same families and is the tested twin.
package metrics

import (
	"strconv"

	bmk "github.com/uppercaveman/k8s-gpu-device-plugin/benchmark"

	"github.com/prometheus/client_golang/prometheus"
)

var labels = []string{"uuid", "index"}

func gauge(name, help string) *prometheus.GaugeVec {
	g := prometheus.NewGaugeVec(prometheus.GaugeOpts{Name: name, Help: help}, labels)
	prometheus.MustRegister(g)
	return g
}

var (
	Healthy    = gauge("gpu_probe_healthy", "1 if every enabled probe passed on the last run")
	Duration   = gauge("gpu_probe_duration_seconds", "Wall time of the last gp_probe_device call")
	CopyGBps   = gauge("gpu_probe_hbm_copy_gbps", "HBM copy bandwidth, algorithmic bytes 2N, median")
	TriadGBps  = gauge("gpu_probe_hbm_triad_gbps", "HBM triad bandwidth, algorithmic bytes 3N, median")
	FillGBps   = gauge("gpu_probe_hbm_fill_gbps", "Pattern fill bandwidth, algorithmic bytes N, median")
	VerifyGBps = gauge("gpu_probe_hbm_verify_gbps", "Pattern verify bandwidth, algorithmic bytes N, median")
	Mismatches = gauge("gpu_probe_pattern_mismatches", "Words that differed from the expected pattern on the last run")
	GemmTFlops = gauge("gpu_probe_gemm_tflops", "bf16 GEMM throughput on tcgen05, median")
	GemmErr    = gauge("gpu_probe_gemm_max_scaled_err", "max |c-ref| / sum|a||b| over the sampled entries")
)

var (
	P2PGBps      = prometheus.NewGaugeVec(prometheus.GaugeOpts{Name: "gpu_probe_p2p_gbps", Help: "NVLink peer bandwidth measured with in-kernel loads/stores"}, []string{"src", "dst", "dir"})
	AllreduceBus = prometheus.NewGaugeVec(prometheus.GaugeOpts{Name: "gpu_probe_allreduce_busbw_gbps", Help: "ncclAllReduce bus bandwidth"}, []string{"ranks"})
	FabricOK     = prometheus.NewGaugeVec(prometheus.GaugeOpts{Name: "gpu_probe_fabric_healthy", Help: "1 if the P2P matrix and the all-reduce met their floors and every transfer carried the right bytes"}, []string{"gpus"})
	FabricSpread = prometheus.NewGaugeVec(prometheus.GaugeOpts{Name: "gpu_probe_fabric_spread", Help: "max/min over the off-diagonal P2P pairs (1.0 = flat)"}, []string{"gpus"})
)

func init() { prometheus.MustRegister(P2PGBps, AllreduceBus, FabricOK, FabricSpread) }

// UpdateFabric is called once per loadPlugins, after gp_probe_fabric.  Twin: PluginManager::GradeFabric.
func UpdateFabric(f bmk.FabricResult) {
	n := strconv.Itoa(f.N)
	for i := 0; i < f.N; i++ {
		for j := 0; j < f.N; j++ {
			if i == j {
				continue
			}
			P2PGBps.WithLabelValues(strconv.Itoa(i), strconv.Itoa(j), "read").Set(f.Read[i*f.N+j])
			P2PGBps.WithLabelValues(strconv.Itoa(i), strconv.Itoa(j), "write").Set(f.Write[i*f.N+j])
		}
	}
	if f.AllreduceError == 0 {
		AllreduceBus.WithLabelValues(n).Set(f.AllreduceBusbwGBps)
	}
	ok := 0.0
	if f.Healthy {
		ok = 1
	}
	FabricOK.WithLabelValues(n).Set(ok)
	FabricSpread.WithLabelValues(n).Set(f.Spread)
}

// Update is called after every probe: start-up grade, periodic tick, GET /probe.
func Update(index string, r bmk.ProbeResult) {
	l := prometheus.Labels{"uuid": r.UUID, "index": index}
	b := 0.0
	if r.Healthy {
		b = 1
	}
	Healthy.With(l).Set(b)
	Duration.With(l).Set(r.DurationSeconds)
	CopyGBps.With(l).Set(r.CopyGBps)
	TriadGBps.With(l).Set(r.TriadGBps)
	FillGBps.With(l).Set(r.FillGBps)
	VerifyGBps.With(l).Set(r.VerifyGBps)
	Mismatches.With(l).Set(float64(r.PatternMismatches))
	GemmTFlops.With(l).Set(r.GemmTFlops)
	GemmErr.With(l).Set(r.GemmMaxScaledErr)
}
