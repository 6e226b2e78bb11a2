// NOT COMPILED in this repository: the build image has no Go toolchain (`go version`: not
// found).  This is the cgo wrapper a maintainer drops into the reference as
// benchmark/probe.go, beside the untouched pprof wrapper (benchmark/benchmark.go:14-124).
// The same calls, in the same order and threading, are exercised from C++ by
// host/plugin_sim.cpp (GpuProbeBackend, host/probe.hpp) and from Python by tests/ through the
// identical C ABI.  INTEGRATION.md §5 maps every function here to its tested C++ twin.
package benchmark

/*
#cgo CFLAGS: -I${SRCDIR}/../include
#cgo LDFLAGS: -L${SRCDIR}/../lib -lgpuprobe
#include <stdlib.h>
#include "gpuprobe.h"
*/
import "C"

import (
	"context"
	"fmt"
	"os"
	"os/exec"
	"runtime"
	"strconv"
	"sync"
	"time"
	"unsafe"
)

const (
	ProbeCopy    = uint32(C.GP_PROBE_COPY)
	ProbeTriad   = uint32(C.GP_PROBE_TRIAD)
	ProbePattern = uint32(C.GP_PROBE_PATTERN)
	ProbeGemm    = uint32(C.GP_PROBE_GEMM)
	ProbeAll     = uint32(C.GP_PROBE_ALL_SINGLE)

	ErrNoMem      = int(C.GP_ERR_NOMEM)
	ErrDeviceLost = int(C.GP_ERR_DEVICE_LOST)
	ErrNccl       = int(C.GP_ERR_NCCL)
)

// Options carries every optional `probe:` key of config.yml (INTEGRATION.md §2); the zero value
// of each field means "library default", exactly as in gp_config.  Twin: ProbeConfig::ToGpConfig
// (host/config.hpp).
type Options struct {
	HbmBytes              uint64
	PatternBytes          uint64 // ^uint64(0) = every free byte (patternBytes: all)
	WoPasses              uint32
	GemmMNK               uint32
	GemmOutputBF16        bool
	MinCopyGBps           float64
	MinTriadGBps          float64
	MinGemmTFlops         float64
	MinP2PGBps            float64
	MaxP2PSpread          float64
	MinAllreduceBusbwGBps float64
	FabricBytes           uint64
	TriadExpectSum        uint64        // known answer of the triad at HbmBytes (INTEGRATION.md §2); 0, 0 = not compared
	TriadExpectXor        uint64
	KeepArena             bool          // keep an idle device's arena between re-probes; Allocate trims it
	SliceBinary           string        // path of gpuprobe-slice; "" disables MIG / isolated probes
	ChildTimeout          time.Duration // 0 -> 120 s
}

// ProbeResult is the Go view of gp_device_result (copied out of C memory).
type ProbeResult struct {
	UUID              string
	Healthy           bool
	Error             int
	FailedMask        uint32
	ProbedMask        uint32
	DurationSeconds   float64
	AllocSeconds      float64
	FreeSeconds       float64
	KernelSeconds     float64
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

// FabricResult is the Go view of gp_fabric_result: Read/Write are n*n, row = source ordinal.
type FabricResult struct {
	N                  int
	UUIDs              []string // by CUDA ordinal
	Read, Write        []float64
	Spread             float64
	WorstSrc, WorstDst int
	WorstGBps          float64
	AllreduceBusbwGBps float64
	AllreduceError     int
	DataOK, Healthy    bool
	Failed             uint32
}

// Probe owns one gp_ctx; create it where plugin/manager.go:44 creates the NVML handle.
type Probe struct {
	mu   sync.Mutex
	ctx  *C.gp_ctx
	opts Options
}

func NewProbe(o Options) (*Probe, error) {
	if uint64(C.gp_sizeof(4)) != uint64(unsafe.Sizeof(C.gp_device_result{})) ||
		uint64(C.gp_sizeof(0)) != uint64(unsafe.Sizeof(C.gp_config{})) ||
		uint64(C.gp_sizeof(7)) != uint64(unsafe.Sizeof(C.gp_fabric_result{})) ||
		uint32(C.gp_abi_version()) != uint32(C.GP_ABI_VERSION) {
		return nil, fmt.Errorf("gpuprobe: struct layout / ABI skew between header and library")
	}
	var cfg C.gp_config
	cfg.struct_size = C.uint32_t(unsafe.Sizeof(cfg))
	cfg.abi_version = C.GP_ABI_VERSION
	cfg.hbm_bytes = C.uint64_t(o.HbmBytes)
	cfg.pattern_bytes = C.uint64_t(o.PatternBytes)
	cfg.wo_passes = C.uint32_t(o.WoPasses)
	cfg.gemm_mnk = C.uint32_t(o.GemmMNK)
	if o.GemmOutputBF16 {
		cfg.gemm_output = 1
	}
	if o.KeepArena {
		cfg.keep_arena = 1
	}
	cfg.min_copy_gbps = C.double(o.MinCopyGBps)
	cfg.min_triad_gbps = C.double(o.MinTriadGBps)
	cfg.min_gemm_tflops = C.double(o.MinGemmTFlops)
	cfg.min_p2p_gbps = C.double(o.MinP2PGBps)
	cfg.max_p2p_spread = C.double(o.MaxP2PSpread)
	cfg.min_allreduce_busbw_gbps = C.double(o.MinAllreduceBusbwGBps)
	cfg.fabric_bytes = C.uint64_t(o.FabricBytes)
	cfg.triad_expect_sum = C.uint64_t(o.TriadExpectSum)
	cfg.triad_expect_xor = C.uint64_t(o.TriadExpectXor)
	var ctx *C.gp_ctx
	// gp_last_error_detail() is thread-local in the library: the goroutine must stay on the OS thread
	// that made the failing call until the text has been fetched.
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if rc := C.gp_init(&cfg, &ctx); rc != 0 {
		return nil, fmt.Errorf("gp_init: %s (%s)", C.GoString(C.gp_strerror(rc)), C.GoString(C.gp_last_error_detail()))
	}
	if o.ChildTimeout == 0 {
		o.ChildTimeout = 120 * time.Second
	}
	return &Probe{ctx: ctx, opts: o}, nil
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
		DurationSeconds: float64(r.duration_s), AllocSeconds: float64(r.alloc_s),
		FreeSeconds: float64(r.free_s), KernelSeconds: float64(r.kernels_s),
		CopyGBps:  float64(r.copy_gbps_median),
		TriadGBps: float64(r.triad_gbps_median), FillGBps: float64(r.fill_gbps_median),
		VerifyGBps: float64(r.verify_gbps_median), PatternPasses: uint32(r.pattern_passes),
		PatternMismatches: uint64(r.pattern_mismatches), PatternDigest: uint64(r.pattern_digest),
		GemmTFlops: float64(r.gemm.tflops_median), GemmMaxScaledErr: float64(r.gemm.max_scaled_err),
	}
}

// ProbeDevice blocks; safe to call concurrently for different UUIDs (the library holds one lock
// per device).  No Go pointer is retained by C past the call.  Twin: GpuProbeBackend::ProbeMany.
func (p *Probe) ProbeDevice(uuid string, mask uint32) (ProbeResult, error) {
	cu := C.CString(uuid)
	defer C.free(unsafe.Pointer(cu))
	var out C.gp_device_result
	runtime.LockOSThread() // the error text below is thread-local in the library
	defer runtime.UnlockOSThread()
	rc := C.gp_probe_device(p.ctx, cu, C.uint32_t(mask), &out)
	res := fromC(&out)
	if rc != 0 {
		if res.Error == 0 {
			res.Error = int(rc)
		}
		return res, fmt.Errorf("gp_probe_device(%s): %s", uuid, C.GoString(C.gp_last_error_detail()))
	}
	return res, nil
}

// ProbeAll grades the first n CUDA devices concurrently (one host thread per GPU inside the library).
func (p *Probe) ProbeAll(n int, mask uint32) ([]ProbeResult, float64, error) {
	out := make([]C.gp_device_result, n)
	var wall C.double
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
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

// ProbeChild grades ONE device from a fresh child process (gpuprobe-slice with
// CUDA_VISIBLE_DEVICES=<uuid>): the only way to reach a MIG slice (device/device.go:32-34), and
// the isolated path for a full GPU whose in-process context was lost.  The child writes the POD
// gp_device_result to its stdout; a child that does not answer before the deadline is killed.
// Twin: GpuProbeBackend::ProbeSlice (posix_spawn + poll deadline + SIGKILL).
func (p *Probe) ProbeChild(uuid string, mask uint32, totalMem uint64) (ProbeResult, error) {
	if p.opts.SliceBinary == "" {
		return ProbeResult{UUID: uuid, Error: int(C.GP_ERR_UNSUPPORTED)}, fmt.Errorf("no slice binary configured")
	}
	hbm := p.opts.HbmBytes
	if hbm == 0 {
		hbm = 16 << 30
	}
	if totalMem != 0 && hbm > totalMem/4 { // triad needs three arrays
		hbm = totalMem / 4
	}
	hbm &^= (1 << 20) - 1
	if hbm < 1<<20 {
		hbm = 1 << 20
	}
	pat := p.opts.PatternBytes
	if pat == 0 || pat > hbm {
		pat = hbm
	}
	ctx, cancel := context.WithTimeout(context.Background(), p.opts.ChildTimeout)
	defer cancel()
	cmd := exec.CommandContext(ctx, p.opts.SliceBinary, strconv.FormatUint(uint64(mask), 10),
		strconv.FormatUint(hbm, 10), strconv.FormatUint(pat, 10), strconv.FormatUint(uint64(p.opts.WoPasses), 10),
		strconv.FormatUint(uint64(p.opts.GemmMNK), 10))
	cmd.Env = append(os.Environ(), "CUDA_VISIBLE_DEVICES="+uuid)
	raw, err := cmd.Output() // CommandContext kills the child at the deadline
	var out C.gp_device_result
	if len(raw) != int(unsafe.Sizeof(out)) {
		return ProbeResult{UUID: uuid, Error: ErrDeviceLost},
			fmt.Errorf("probe process for %s returned no result (%v)", uuid, err)
	}
	copy((*[unsafe.Sizeof(out)]byte)(unsafe.Pointer(&out))[:], raw) // same machine, same layout: the POD struct as written
	res := fromC(&out)
	if res.Error != 0 || !res.Healthy {
		return res, fmt.Errorf("probe process for %s: gp_error %d failed_mask %d", uuid, res.Error, res.FailedMask)
	}
	return res, nil
}

// ProbeFabric measures the P2P matrices and the all-reduce over the first n CUDA devices and
// grades them against the fabric floors.  Twin: GpuProbeBackend::ProbeFabric.
func (p *Probe) ProbeFabric(n int) (FabricResult, error) {
	var out C.gp_fabric_result
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if rc := C.gp_probe_fabric(p.ctx, C.int(n), &out); rc != 0 {
		return FabricResult{}, fmt.Errorf("gp_probe_fabric: %s", C.GoString(C.gp_last_error_detail()))
	}
	f := FabricResult{N: n, Spread: float64(out.spread), WorstSrc: int(out.worst_src), WorstDst: int(out.worst_dst),
		WorstGBps: float64(out.worst_gbps), AllreduceBusbwGBps: float64(out.allreduce_busbw_gbps),
		AllreduceError: int(out.allreduce_error), DataOK: out.data_ok == 1, Healthy: out.healthy == 1,
		Failed: uint32(out.failed)}
	for i := 0; i < n*n; i++ {
		f.Read = append(f.Read, float64(out.read_gbps[i]))
		f.Write = append(f.Write, float64(out.write_gbps[i]))
	}
	buf := (*C.char)(C.malloc(64))
	defer C.free(unsafe.Pointer(buf))
	for i := 0; i < n; i++ {
		C.gp_device_uuid(p.ctx, C.int(i), buf, 64)
		f.UUIDs = append(f.UUIDs, C.GoString(buf))
	}
	return f, nil
}

// OtherComputeProcs: compute processes other than this one on the device, -1 when NVML cannot
// say.  Twin: GpuProbeBackend::OtherComputeProcs.
func (p *Probe) OtherComputeProcs(uuid string) int {
	cu := C.CString(uuid)
	defer C.free(unsafe.Pointer(cu))
	var info C.gp_nvml_info
	if C.gp_nvml_snapshot(p.ctx, cu, &info) != 0 || info.available == 0 {
		return -1
	}
	return int(info.other_compute_procs)
}

// Trim returns every arena to the driver once grading is done (pods need the memory).
func (p *Probe) Trim() { C.gp_trim(p.ctx) }

// TrimDevice returns ONE device's kept arena; called from Allocate before the pod starts.
func (p *Probe) TrimDevice(uuid string) {
	cu := C.CString(uuid)
	defer C.free(unsafe.Pointer(cu))
	C.gp_trim_device(p.ctx, cu)
}

// ResetDevice recovers a context lost to a sticky fault without restarting the plugin
// (cudaDeviceReset; nothing else in the plugin uses CUDA).
func (p *Probe) ResetDevice(uuid string) int {
	cu := C.CString(uuid)
	defer C.free(unsafe.Pointer(cu))
	return int(C.gp_device_reset(p.ctx, cu))
}
