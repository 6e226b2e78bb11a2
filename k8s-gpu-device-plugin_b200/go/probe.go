// NOT COMPILED in this repository: the build image has no Go toolchain (`go version`: not
// found).  This is the cgo wrapper a maintainer drops into the reference as
// benchmark/probe.go, beside the untouched pprof wrapper (benchmark/benchmark.go:14-124).
// The same calls, in the same order and threading, are exercised from C++ by
// host/plugin_sim.cpp and from Python by tests/ through the identical C ABI.
package benchmark

/*
#cgo CFLAGS: -I${SRCDIR}/../include
#cgo LDFLAGS: -L${SRCDIR}/../lib -lgpuprobe
#include <stdlib.h>
#include "gpuprobe.h"
*/
import "C"

import (
	"fmt"
	"sync"
	"unsafe"
)

const (
	ProbeCopy    = uint32(C.GP_PROBE_COPY)
	ProbeTriad   = uint32(C.GP_PROBE_TRIAD)
	ProbePattern = uint32(C.GP_PROBE_PATTERN)
	ProbeGemm    = uint32(C.GP_PROBE_GEMM)
	ProbeAll     = uint32(C.GP_PROBE_ALL_SINGLE)
)

// ProbeResult is the Go view of gp_device_result (copied out of C memory).
type ProbeResult struct {
	UUID              string
	Healthy           bool
	Error             int
	FailedMask        uint32
	ProbedMask        uint32
	DurationSeconds   float64
	CopyGBps          float64
	TriadGBps         float64
	FillGBps          float64
	VerifyGBps        float64
	PatternPasses     uint32
	PatternMismatches uint64
	PatternDigest     uint64
	GemmTFlops        float64
	GemmMaxScaledErr  float64
}

// Probe owns one gp_ctx; create it where plugin/manager.go:44 creates the NVML handle.
type Probe struct {
	mu  sync.Mutex
	ctx *C.gp_ctx
}

func NewProbe(hbmBytes uint64, minCopyGBps float64) (*Probe, error) {
	if uint64(C.gp_sizeof(4)) != uint64(unsafe.Sizeof(C.gp_device_result{})) {
		return nil, fmt.Errorf("gpuprobe: struct layout skew between header and library")
	}
	var cfg C.gp_config
	cfg.struct_size = C.uint32_t(unsafe.Sizeof(cfg))
	cfg.abi_version = C.GP_ABI_VERSION
	cfg.hbm_bytes = C.uint64_t(hbmBytes)
	cfg.min_copy_gbps = C.double(minCopyGBps)
	var ctx *C.gp_ctx
	if rc := C.gp_init(&cfg, &ctx); rc != 0 {
		return nil, fmt.Errorf("gp_init: %s (%s)", C.GoString(C.gp_strerror(rc)), C.GoString(C.gp_last_error_detail()))
	}
	return &Probe{ctx: ctx}, nil
}

func (p *Probe) Close() {
	p.mu.Lock()
	defer p.mu.Unlock()
	if p.ctx != nil {
		C.gp_shutdown(p.ctx)
		p.ctx = nil
	}
}

func fromC(r *C.gp_device_result) ProbeResult {
	return ProbeResult{
		UUID: C.GoString(&r.uuid[0]), Healthy: r.healthy == 1, Error: int(r.error),
		FailedMask: uint32(r.failed_mask), ProbedMask: uint32(r.probed_mask),
		DurationSeconds: float64(r.duration_s), CopyGBps: float64(r.copy_gbps_median),
		TriadGBps: float64(r.triad_gbps_median), FillGBps: float64(r.fill_gbps_median),
		VerifyGBps: float64(r.verify_gbps_median), PatternPasses: uint32(r.pattern_passes),
		PatternMismatches: uint64(r.pattern_mismatches), PatternDigest: uint64(r.pattern_digest),
		GemmTFlops: float64(r.gemm.tflops_median), GemmMaxScaledErr: float64(r.gemm.max_scaled_err),
	}
}

// ProbeDevice blocks; safe to call concurrently for different UUIDs (the library holds one lock
// per device).  No Go pointer is retained by C past the call.
func (p *Probe) ProbeDevice(uuid string, mask uint32) (ProbeResult, error) {
	cu := C.CString(uuid)
	defer C.free(unsafe.Pointer(cu))
	var out C.gp_device_result
	rc := C.gp_probe_device(p.ctx, cu, C.uint32_t(mask), &out)
	res := fromC(&out)
	if rc != 0 {
		return res, fmt.Errorf("gp_probe_device(%s): %s", uuid, C.GoString(C.gp_last_error_detail()))
	}
	return res, nil
}

// ProbeAll grades the first n CUDA devices concurrently (one host thread per GPU inside the library).
func (p *Probe) ProbeAll(n int, mask uint32) ([]ProbeResult, float64, error) {
	out := make([]C.gp_device_result, n)
	var wall C.double
	rc := C.gp_probe_all(p.ctx, C.int(n), C.uint32_t(mask), &out[0], &wall)
	res := make([]ProbeResult, n)
	for i := range out {
		res[i] = fromC(&out[i])
	}
	if rc != 0 {
		return res, float64(wall), fmt.Errorf("gp_probe_all: %s", C.GoString(C.gp_last_error_detail()))
	}
	return res, float64(wall), nil
}

// Trim returns every arena to the driver once grading is done (pods need the memory).
func (p *Probe) Trim() { C.gp_trim(p.ctx) }
