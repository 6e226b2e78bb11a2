#!/usr/bin/env python3
"""bench.py — the HBM probe of BASELINE.json (`metric`: HBM probe GB/s per GPU; full-box
health-sweep seconds), measured through the C ABI of libgpuprobe.so.

    python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, sm_100a)
    python bench.py --impl reference --gpus N --steps K ...  # CPU port of the spec on host cores

One "step" = one pass of the hot path over one batch of synthetic device-resident
input (gp_hbm_step, include/gpuprobe.h): fill X (walking-ones) -> copy X->Y -> verify Y ->
fill X (address-in-address) -> verify X -> triad Z=b+3c -> checksum Z, over 16 GiB per
array (BASELINE.json configs[1]).  Algorithmic bytes per step = 10 * bytes
(fill N, copy 2N, verify N, triad 3N).  Every step's three 48-byte results
are read back to the host and checked against closed forms inside the timed region.

Multi-GPU (torchrun, one rank per GPU): the path shards as independent devices, so
ranks run the same per-GPU workload with no data-path collective ("scaling": "weak");
torch.distributed (NCCL) is used only for the barrier and the max-over-ranks time.

Beside `value` the line carries: `e2e` (gp_probe_device through the C ABI, the periodic
re-probe of an idle device with its arena kept) and `e2e_cold` (the same call allocating and
releasing every time, with alloc/free/kernel seconds); `sweep` (gp_probe_all over all N GPUs
from rank 0: the second half of the metric); `gemm` (configs[2]); and at N >= 2, from rank 0
with the other ranks parked on a host barrier, the only cross-GPU paths (configs[4]): `p2p`
(in-kernel read / write / bidirectional matrices at 1 GiB with every transfer's data checked,
the cudaMemcpyPeerAsync yardstick in the same schedule, the all-at-once round and the 64 KiB -
1 GiB ladder), `allreduce` (NCCL bus bandwidth over the ladder + 4 GiB) and `fabric` (the verdict).

The reference (uppercaveman/k8s-gpu-device-plugin) has no implementation of this path
(benchmark/benchmark.go is a pprof wrapper) and no Go toolchain exists here, so the
`--impl reference` arm and `cpu_baseline` time this repo's CPU port of the frozen spec
(oracle/gpuprobe_oracle.c, kind "port") — the one place bench.py executes oracle/.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-device-plugin_b200"))

M64 = (1 << 64) - 1
GIB = 1 << 30
UNIT = "GB/s"
METRIC = "hbm_probe_gbps"


def closed_forms(nwords, step):
    """Expected (sum, xor) of the step's two pattern verifies — SURVEY.md §8(c) closed forms."""
    wo = ((-(nwords // 64)) & M64, 0 if (nwords // 64) % 2 == 0 else M64)
    big_x = lambda m: [m, 1, m + 1, 0][m % 4]
    s, x = (4 * nwords * (nwords - 1)) & M64, (8 * big_x(nwords - 1)) & M64
    if step & 1:
        s, x = (nwords * M64 - s) & M64, x ^ (M64 if nwords % 2 else 0)
    return wo, (s, x)


def measured_gemm_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            j = json.load(f)
        return float(j["bf16_tflops"]), float(j.get("bf16_tflops_sustained", 0) or 0), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 1590.0, 1400.0, "fallback (B200_PROFILING.md 1.59 PFLOP/s burst)"


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            v = float(json.load(f)["hbm_gbs"])
        return v, "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML)."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
               0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, index, period=0.05, uuid=None):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.stop_ev = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            # CUDA and NVML may enumerate in different orders: address the GPU by UUID when we have it
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if isinstance(uuid, str) else uuid) if uuid else None
            except Exception:
                self.h = None
            if self.h is None:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # no NVML: report that, do not invent clocks
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        while not self.stop_ev.is_set():
            try:
                self.samples.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                try:
                    mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self.stop_ev.wait(self.period)

    def finish(self):
        self.stop_ev.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def max_over_ranks(dist, values, device):
    """MAX over ranks of a list of floats (device-timed intervals); identity when not distributed."""
    if dist is None:
        return list(values)
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t]


def dist_env():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


# ------------------------------------------------------------------ CPU port ----------------
def cpu_port_run(nbytes, steps, warmup, min_seconds=0.0, max_seconds=30.0, threads=None):
    """Times the CPU port of the step (oracle/) with every host thread (or `threads`).  Returns a dict."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib
    lib = oracle_lib.load()
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    if threads:
        cores = threads
    arena = np.zeros(5 * nbytes, dtype=np.uint8)  # first touch here
    nwords = nbytes // 8
    for w in range(max(warmup, 1)):
        oracle_lib.hbm_step(lib, arena, nbytes, step=w, gen=(w == 0), nthreads=cores)
    done, t0 = 0, time.perf_counter()
    while True:
        out = oracle_lib.hbm_step(lib, arena, nbytes, step=done, gen=False, nthreads=cores)
        wo, ad = closed_forms(nwords, done)
        assert (out[0].sum, out[0].xor_, out[0].mismatches) == (wo[0], wo[1], 0)
        assert (out[1].sum, out[1].xor_, out[1].mismatches) == (ad[0], ad[1], 0)
        done += 1
        el = time.perf_counter() - t0
        if (done >= steps and el >= min_seconds) or el >= max_seconds:
            break
    el = time.perf_counter() - t0
    return {"value": 10.0 * nbytes * done / el / 1e9, "seconds": el, "steps": done, "cores": cores,
            "ms_per_step": el / done * 1e3,
            "sample": f"{done} steps of the same 7-op step over {nbytes >> 20} MiB per array "
                      f"(BASELINE.json configs[0] size) on {cores} threads"}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    nbytes = args.cpu_bytes
    r = cpu_port_run(nbytes, args.steps, max(args.warmup, 1), max_seconds=120.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": r["steps"], "warmup": max(args.warmup, 1), "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64+f32",
        "data": "synthetic",
        "config": {"workload": "1xB200 HBM stream probe: copy/triad + bit-exact pattern verify "
                               "(BASELINE.json configs[1]); bounded CPU sample at configs[0] size",
                   "bytes_per_array": nbytes, "arrays": 5, "algorithmic_bytes_per_step": 10 * nbytes,
                   "host": "B200 host cores, no GPU"},
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                         "sample": r["sample"],
                         "note": "CPU port of the frozen spec; the reference has no code for this path"},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------ K4 leg ------------------
LADDER = [64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30]
NVLINK_GBPS = 900.0  # per direction per GPU


def _offdiag(m):
    n = len(m)
    return [m[i][j] for i in range(n) for j in range(n) if i != j]


def _summ(m):
    off = _offdiag(m)
    n = len(m)
    return {"min": round(min(off), 1), "max": round(max(off), 1), "mean": round(sum(off) / len(off), 1),
            "frac_of_900": round(sum(off) / len(off) / NVLINK_GBPS, 4), "min_frac_of_900": round(min(off) / NVLINK_GBPS, 4),
            "flatness_max_over_min": round(max(off) / min(off), 4),
            "local_mean": round(sum(m[i][i] for i in range(n)) / n, 1)}


def fabric_leg(gpuprobe, n, nbytes):
    """P2P matrices (in-kernel peer loads / stores, 1-factorisation rounds), the copy-engine yardstick in the
    same schedule, the 64 KiB - 1 GiB ladder, the NCCL all-reduce ladder and the fabric verdict.  Every read
    is checked against the closed form of what the peer holds, every write / copy where it landed."""
    out = {}
    t0 = time.perf_counter()
    try:
        with gpuprobe.Probe(keep_arena=1, hbm_bytes=GIB, arena_cap_bytes=9 * GIB, fabric_bytes=nbytes) as p:
            modes = [("read", gpuprobe.GP_P2P_READ), ("write", gpuprobe.GP_P2P_WRITE),
                     ("bidir_write", gpuprobe.GP_P2P_BIDIR_WRITE), ("bidir_read", gpuprobe.GP_P2P_BIDIR_READ),
                     ("copy_engine", gpuprobe.GP_P2P_COPY_ENGINE),
                     ("copy_engine_bidir", gpuprobe.GP_P2P_COPY_ENGINE_BIDIR)]
            p2p = {"bytes": nbytes, "n_gpus": n, "unit": UNIT, "peak": NVLINK_GBPS,
                   "what": "GB/s per direction per ordered pair, best of 3 after a warm-up, CUDA events on the "
                           "source GPU's stream; pairs scheduled as the n-1 rounds of a 1-factorisation; "
                           "copy_engine* = cudaMemcpyPeerAsync in the same schedule (the yardstick)"}
            for name, d in modes:
                p2p[name] = _summ(p.probe_p2p(n, nbytes, d))
            m = p.probe_p2p(n, nbytes, gpuprobe.GP_P2P_ALL_AT_ONCE)
            agg = [m[i][i] for i in range(n)]
            p2p["all_at_once"] = {"aggregate_ingress_per_gpu": {"min": round(min(agg), 1), "max": round(max(agg), 1),
                                                                "mean": round(sum(agg) / n, 1)},
                                  "frac_of_900": round(sum(agg) / n / NVLINK_GBPS, 4)}
            p2p["write_vs_copy_engine"] = round(p2p["write"]["mean"] / p2p["copy_engine"]["mean"], 4)
            p2p["read_vs_copy_engine"] = round(p2p["read"]["mean"] / p2p["copy_engine"]["mean"], 4)
            ladder = {}
            for name, d in (("read", gpuprobe.GP_P2P_READ), ("write", gpuprobe.GP_P2P_WRITE),
                            ("copy_engine", gpuprobe.GP_P2P_COPY_ENGINE)):
                ladder[name] = {str(sz): round(sum(_offdiag(mm)) / (n * (n - 1)), 1)
                                for sz in LADDER for mm in [p.probe_p2p(n, sz, d)]}
            p2p["ladder_mean_gbps"] = ladder
            out["p2p"] = p2p
            f = p.probe_fabric(n)
            out["fabric"] = {"healthy": int(f.healthy), "data_ok": int(f.data_ok), "failed_bits": int(f.failed),
                             "spread": round(f.spread, 4), "worst_pair": [f.worst_src, f.worst_dst],
                             "worst_gbps": round(f.worst_gbps, 1), "seconds": round(f.duration_s, 3),
                             "allreduce_busbw_gbps": round(f.allreduce_busbw_gbps, 1)}
            ar = {"nranks": n, "unit": UNIT, "what": "in-place fp32 sum ncclAllReduce, one communicator per GPU in "
                  "one process, group launch, max over ranks of the device time; busbw = algbw * 2(n-1)/n"}
            bus = {}
            for sz in LADDER + [4 << 30]:
                try:
                    a, b = p.probe_allreduce(n, sz, warmup=2, iters=5)
                    bus[str(sz)] = round(b, 1)
                except gpuprobe.GpuProbeError as e:
                    bus[str(sz)] = f"error: {e}"
                    break
            ar["busbw_gbps"] = bus
            good = [v for v in bus.values() if isinstance(v, float)]
            ar["busbw_peak_gbps"] = max(good) if good else None
            # n = 2, 4, 8 — the smaller rank counts at the two largest sizes
            sub = {}
            for nr in (2, 4):
                if nr < n:
                    try:
                        sub[str(nr)] = {str(sz): round(p.probe_allreduce(nr, sz, warmup=2, iters=5)[1], 1)
                                        for sz in (1 << 30, 4 << 30)}
                    except gpuprobe.GpuProbeError as e:
                        sub[str(nr)] = f"error: {e}"
            if sub:
                ar["busbw_gbps_by_ranks"] = sub
            out["allreduce"] = ar
    except gpuprobe.GpuProbeError as e:
        out.setdefault("p2p", {})["error"] = str(e)
    out.setdefault("p2p", {})["leg_seconds"] = round(time.perf_counter() - t0, 2)
    return out


# ------------------------------------------------------------------ CUDA arm ----------------
def run_ours(args):
    import gpuprobe
    rank, local_rank, world = dist_env()
    n_gpus = args.gpus
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the probe has no CPU fallback "
                         "(use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    # a host-side barrier for the sweep leg: an NCCL barrier would park a spinning kernel on
    # every waiting rank's GPU while rank 0 grades those same GPUs
    cpu_group = dist.new_group(backend="gloo") if dist is not None else None

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def host_barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier(group=cpu_group)

    nbytes = args.bytes
    nwords = nbytes // 8
    variant = {"auto": 0, "vec": 1, "tma": 2}[args.variant]
    probe = gpuprobe.Probe(keep_arena=1, hbm_bytes=nbytes, arena_cap_bytes=5 * nbytes + GIB,
                           stream_variant=variant)
    dev = probe.device_uuid(local_rank)

    def check(out, step):
        wo, ad = closed_forms(nwords, step)
        if (out[0].sum, out[0].xor_, out[0].mismatches) != (wo[0], wo[1], 0):
            raise SystemExit(f"step {step}: walking-ones verify of the copy destination failed: {out[0].astuple()}")
        if (out[1].sum, out[1].xor_, out[1].mismatches) != (ad[0], ad[1], 0):
            raise SystemExit(f"step {step}: address verify failed: {out[1].astuple()}")

    # the GEMM probe is a kernel timed alone (compared with the measured BURST cuBLAS figure), so it
    # runs first, on a GPU that has not yet been heated by the HBM steps
    gemm = None
    if rank == 0 and not args.no_sweep:
        try:
            g = probe.gemm_probe(dev, mnk=8192, iters=10, warmup=2, samples=4096)
            burst, sustained, src = measured_gemm_peaks()
            gemm = {"workload": "bf16 GEMM probe M=N=K=8192 on tcgen05/TMEM (BASELINE.json configs[2])",
                    "tflops_best": g.tflops_best, "tflops_median": g.tflops_median, "best_ms": g.t.best_ms,
                    "bound": "tensor", "peak": burst, "peak_source": src, "frac": g.tflops_best / burst,
                    "frac_of_sustained_cublas": (g.tflops_median / sustained) if sustained else None,
                    "frac_of_nominal_2250": g.tflops_best / 2250.0,
                    "max_scaled_err": g.max_scaled_err, "freivalds_err": g.freivalds_err, "tolerance": 1e-3,
                    "samples": g.samples, "samples_bad": g.samples_bad}
            try:
                with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                    gj = json.load(f).get("gemm")
                if gj:
                    gemm["ncu"] = gj   # tensor-pipe active %, DRAM bytes per launch, source file (a profiler capture, not this run)
            except Exception:
                pass
            # the same problem with C in bf16, which is what the measured cuBLAS bf16 peak (torch.matmul of
            # bf16 tensors) writes; the bf16 C must equal RNE(bf16) of the fp32 C of the same plan bit for bit
            with gpuprobe.Probe(keep_arena=1, gemm_output=1) as pg:
                gb = pg.gemm_probe(dev, mnk=8192, iters=10, warmup=2, samples=4096)
            gemm["bf16_output"] = {"tflops_best": gb.tflops_best, "tflops_median": gb.tflops_median,
                                   "frac": gb.tflops_best / burst, "frac_of_nominal_2250": gb.tflops_best / 2250.0,
                                   "max_scaled_err": gb.max_scaled_err, "freivalds_err": gb.freivalds_err,
                                   "samples_bad": gb.samples_bad, "bf16_mismatches_vs_rne_fp32": gb.bf16_mismatches}
        except gpuprobe.GpuProbeError as e:
            gemm = {"error": str(e)} if gemm is None else dict(gemm, bf16_output={"error": str(e)})

    host_barrier()
    triad_sig = None
    for w in range(args.warmup):
        out = probe.hbm_step(dev, nbytes, w)
        check(out, w)
        triad_sig = (out[2].sum, out[2].xor_, out[2].wsum)

    BAD = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    attempts = 0
    while True:
        attempts += 1
        probe.kernel_stats(dev, reset=True)
        sampler = ClockSampler(local_rank, uuid=dev)
        sampler.start()
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            out = probe.hbm_step(dev, nbytes, k)   # blocks; 144 B of results come back D2H each step
            check(out, k)
            if triad_sig is not None and (out[2].sum, out[2].xor_, out[2].wsum) != triad_sig:
                raise SystemExit(f"step {k}: triad checksum changed between steps")
        barrier()
        wall = time.perf_counter() - t0
        clocks = sampler.finish()
        # a run that saw a hardware / thermal slowdown is taken again, once
        throttled = bool(BAD & set(clocks.get("reasons", [])))
        if max_over_ranks(dist, [1.0 if throttled else 0.0], "cuda")[0] == 0.0 or attempts >= 2:
            break
    clocks["attempts"] = attempts
    stats = probe.kernel_stats(dev)
    dev_ms = stats[gpuprobe.GP_K_STEP].total_ms

    wall_max, dev_ms_max = max_over_ranks(dist, [wall, dev_ms], "cuda")

    e2e, e2e_cold = None, None
    probe.close()
    mask = gpuprobe.GP_PROBE_COPY | gpuprobe.GP_PROBE_TRIAD | gpuprobe.GP_PROBE_PATTERN

    def e2e_leg(p, calls, what):
        """`calls` back-to-back gp_probe_device calls on every rank at once, host clock, max over ranks."""
        p.kernel_stats(dev, reset=True)
        barrier()
        t0 = time.perf_counter()
        res = []
        for _ in range(calls):
            r = p.probe_device(dev, mask)   # generate, warm up, time, 64+2 fill/verify passes, verify, results D2H
            if not r.healthy:
                raise SystemExit(f"e2e: gp_probe_device verdict unhealthy, failed_mask={r.failed_mask}")
            res.append(r)
        barrier()
        w = time.perf_counter() - t0
        st2 = p.kernel_stats(dev)
        nb = sum(st2[i].algo_bytes for i in (gpuprobe.GP_K_COPY, gpuprobe.GP_K_TRIAD,
                                             gpuprobe.GP_K_FILL, gpuprobe.GP_K_VERIFY))
        launches = sum(int(st2[i].launches) for i in range(7))
        mean = lambda f: sum(getattr(r, f) for r in res) / len(res)
        w, a_s, f_s, k_s = max_over_ranks(dist, [w, mean("alloc_s"), mean("free_s"), mean("kernels_s")], "cuda")
        d2h = ctypes.sizeof(gpuprobe.DeviceResult) + res[-1].pattern_passes * 48 + 48 + 8
        return {"value": nb * n_gpus / w / 1e9, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": d2h,
                "calls": calls, "seconds_per_call": w / calls, "launches_per_call": launches // calls,
                "alloc_s_per_call": a_s, "free_s_per_call": f_s, "kernels_s_per_call": k_s,
                "arena_bytes": int(res[-1].arena_bytes),
                "arena_backend": {1: "vmm", 2: "cudaMalloc"}.get(int(res[-1].arena_backend), "?"),
                "arena_reused": int(res[-1].arena_reused), "what": what}

    if args.no_e2e:
        p2 = gpuprobe.Probe(keep_arena=0, hbm_bytes=nbytes, stream_variant=variant)
    else:
        e2e_calls = max(1, min(args.steps, args.e2e_calls))
        # ---- e2e: the call the plugin makes on every periodic re-probe of an idle device.  The arena
        # stays mapped while the device is unassigned (keep_arena = 1) and goes back to the pods in
        # Allocate (gp_trim_device), so the steady-state call allocates nothing.
        pk = gpuprobe.Probe(keep_arena=1, hbm_bytes=nbytes, stream_variant=variant)
        pk.probe_device(dev, mask)  # first call of the process: context, module load, arena
        e2e = e2e_leg(pk, e2e_calls,
                      "gp_probe_device(uuid, COPY|TRIAD|PATTERN) through the C ABI, arena kept between calls "
                      "(keep_arena = 1: the periodic re-probe of an idle device; gp_trim_device returns the "
                      "memory in Allocate): operand generation, warm-up + timed launches, 64+2 fill/verify "
                      "passes, verification, results copied to the caller's host struct.  The probe takes no "
                      "bulk host input by construction (patterns are generated on the device it grades), so "
                      "h2d is 0 and d2h is the result bytes.")
        pk.close()
        # ---- the same call with a cold arena every time (keep_arena = 0: the start-up grade); alloc_s and
        # free_s say what the allocator costs when all N processes hit the driver at once
        p2 = gpuprobe.Probe(keep_arena=0, hbm_bytes=nbytes, stream_variant=variant)
        p2.probe_device(dev, mask)
        e2e_cold = e2e_leg(p2, e2e_calls, "same call with keep_arena = 0, back to back: cudaMalloc at entry, the arena "
                                          "handed to a release thread at return; each call first waits for its "
                                          "predecessor's release (alloc_s_per_call is that wait plus the allocation)")

    # ---- the literal host-buffer reading of e2e, for the record: grade 1 GiB that lives in pinned HOST
    # memory (H2D copy + verify + 48-byte result back).  This measures PCIe, not the device, and is not
    # what the probe is for — the probe grades device memory with patterns generated on the device.
    e2e_host = None
    if not args.no_e2e:
        try:
            import numpy as np
            hb_words = (1 << 30) // 8
            pinned = torch.empty(hb_words, dtype=torch.int64).pin_memory()
            host_np = pinned.numpy().view(np.uint64)
            idx = np.arange(hb_words, dtype=np.uint64)
            host_np[:] = np.left_shift(np.uint64(1), (idx + np.uint64(7)) & np.uint64(63))
            del idx
            p2.pattern_host(dev, host_np, gpuprobe.GP_PAT_WALKING_ONES, 7, fill=False)  # warm-up, arena alloc
            times = []
            for _ in range(3):
                t0 = time.perf_counter()
                r = p2.pattern_host(dev, host_np, gpuprobe.GP_PAT_WALKING_ONES, 7, fill=False)
                times.append(time.perf_counter() - t0)
                if r.mismatches:
                    raise SystemExit("host-buffer verify found mismatches")
            best = min(times)
            e2e_host = {"value": (1 << 30) / best / 1e9, "unit": UNIT, "h2d_bytes_per_step": 1 << 30,
                        "d2h_bytes_per_step": 48, "seconds_per_call": best,
                        "what": "gp_pattern_host(fill=0): verify of a 1 GiB pinned HOST buffer through the C ABI; "
                                "PCIe-bound by construction, reported for completeness, not the product path"}
            p2.trim()
        except gpuprobe.GpuProbeError as e:
            e2e_host = {"error": str(e)}

    # ---- full-box sweep seconds (second half of the metric), rank 0, all N GPUs, one process --
    sweep = None
    if not args.no_sweep:
        p2.trim()
        host_barrier()
        if rank == 0:
            p3 = gpuprobe.Probe(keep_arena=0, hbm_bytes=nbytes, stream_variant=variant)
            def timed_sweep(mask):
                # first call of the process on these GPUs (context creation, module load), then the memory of that
                # call is allowed to arrive back at the driver (gp_trim joins the release threads) before the
                # timed call starts: a cold-arena sweep as the plugin runs it, not one queued behind its predecessor
                res, first = p3.probe_all(n_gpus, mask)
                p3.trim()
                res, wall = p3.probe_all(n_gpus, mask)
                t0 = time.perf_counter()
                p3.trim()
                return res, first, wall, time.perf_counter() - t0
            try:
                res, wall_first, wall, rel = timed_sweep(gpuprobe.GP_PROBE_ALL_SINGLE)
                sweep = {"n_gpus": n_gpus, "seconds": wall, "seconds_first_call": wall_first,
                         "release_after_return_s": rel,
                         "alloc_wait_s": [round(r.alloc_s, 4) for r in res],
                         "healthy": [int(r.healthy) for r in res],
                         "copy_gbps": [round(r.copy_gbps_median, 1) for r in res],
                         "triad_gbps": [round(r.triad_gbps_median, 1) for r in res],
                         "fill_gbps": [round(r.fill_gbps_median, 1) for r in res],
                         "verify_gbps": [round(r.verify_gbps_median, 1) for r in res],
                         "gemm_tflops": [round(r.gemm.tflops_median, 1) for r in res],
                         "gemm_max_scaled_err": [r.gemm.max_scaled_err for r in res],
                         "probes": "copy+triad+pattern(64+2 passes)+gemm, 16 GiB/array, cold arena (keep_arena = 0): "
                                   "gp_probe_all entry to return incl. allocation and result copy-out; the arenas are "
                                   "released on background threads as the call returns (release_after_return_s)"}
            except gpuprobe.GpuProbeError as e:
                try:
                    mask3 = gpuprobe.GP_PROBE_COPY | gpuprobe.GP_PROBE_TRIAD | gpuprobe.GP_PROBE_PATTERN
                    res, wall_first, wall, rel = timed_sweep(mask3)
                    sweep = {"n_gpus": n_gpus, "seconds": wall, "seconds_first_call": wall_first,
                             "release_after_return_s": rel,
                             "healthy": [int(r.healthy) for r in res], "probes": "copy+triad+pattern",
                             "note": f"gemm probe unavailable: {e}"}
                except gpuprobe.GpuProbeError as e2:
                    sweep = {"error": str(e2)}
            p3.close()
        host_barrier()
    # ---- K4: the only cross-GPU paths, rank 0 drives all N GPUs from one process (the plugin's shape)
    fabric = None
    if n_gpus >= 2 and not args.no_fabric:
        p2.trim()
        host_barrier()
        if rank == 0:
            fabric = fabric_leg(gpuprobe, n_gpus, args.fabric_bytes)
        host_barrier()
    host_barrier()
    p2.close()

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peak()
    step_bytes = 10.0 * nbytes
    value = step_bytes * args.steps * n_gpus / wall_max / 1e9
    kernels = {}
    for i in (gpuprobe.GP_K_COPY, gpuprobe.GP_K_TRIAD, gpuprobe.GP_K_FILL, gpuprobe.GP_K_VERIFY):
        s = stats[i]
        if s.launches:
            ach = s.algo_bytes / (s.total_ms * 1e-3) / 1e9
            kernels[gpuprobe.GP_K_NAMES[i]] = {
                "launches": int(s.launches), "avg_ms": s.total_ms / s.launches,
                "achieved_gbps": ach, "frac_of_measured_peak": ach / peak, "frac_of_8tbps": ach / 8000.0,
                "share_of_step": s.total_ms / dev_ms if dev_ms else None}
    dom = max(kernels, key=lambda k: kernels[k]["share_of_step"])
    traffic, traffic_note = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f).get(dom)
        alg = {"copy": 2, "triad": 3, "fill": 1, "verify": 1}[dom] * nbytes
        key = f"dram_bytes_per_launch_at_{nbytes}"
        if tj and key in tj:
            traffic, traffic_note = tj[key], tj.get("source_full_size")
        elif tj:
            traffic = tj["ratio_traffic_over_algorithmic"] * alg
            traffic_note = ("ncu capture at 2 GiB per array scaled by size: measured traffic/algorithmic = "
                            f"{tj['ratio_traffic_over_algorithmic']:.4f}; " + tj.get("source", ""))
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["achieved_gbps"], "peak": peak,
                "unit": "GB/s", "frac": kernels[dom]["achieved_gbps"] / peak, "traffic": traffic,
                "traffic_note": traffic_note,
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": {"copy": 2 * nbytes, "triad": 3 * nbytes,
                                                 "fill": nbytes, "verify": nbytes}[dom]}
    # every verify is two kernels (chunk grid + one-CTA finalize): 10 launches per step
    gpu_launches = sum(int(stats[i].launches) for i in range(7)) + int(stats[gpuprobe.GP_K_VERIFY].launches)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": wall_max / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64+f32", "data": "synthetic",
        "config": {"workload": "1xB200 HBM stream probe: 16 GiB copy/triad + bit-exact pattern verify "
                               "(BASELINE.json configs[1]), per GPU",
                   "bytes_per_array": nbytes, "arrays": 5, "algorithmic_bytes_per_step": int(step_bytes),
                   "step": "fill WO, copy, verify, fill AiA, verify, triad, checksum",
                   "stream_variant": args.variant, "parallelism": f"replicas x{n_gpus} (no collective)",
                   "l2": "inputs (16 GiB per array) are larger than the 126 MB L2; no flush needed"},
        "per_gpu_gbps": value / n_gpus, "frac_of_8tbps_per_gpu": value / n_gpus / 8000.0,
        "device_ms_per_step": dev_ms_max / args.steps,
        "roofline": roofline, "kernels": kernels, "clocks": clocks,
        "e2e": e2e,
        "gpu_launches": gpu_launches,
    }
    if e2e_cold is not None:
        line["e2e_cold"] = e2e_cold
    if fabric is not None:
        line.update(fabric)
    if sweep is not None:
        line["sweep"] = sweep
    if gemm is not None:
        line["gemm"] = gemm
    if e2e_host is not None:
        line["e2e_host_buffer"] = e2e_host
    if n_gpus == 1 and not args.no_cpu:
        c = cpu_port_run(args.cpu_bytes, 1, 1, min_seconds=10.0, max_seconds=30.0)
        c1 = cpu_port_run(args.cpu_bytes, 1, 1, min_seconds=3.0, max_seconds=8.0, threads=1)
        line["cpu_baseline"] = {"value": c["value"], "unit": UNIT, "cores": c["cores"], "kind": "port",
                                "sample": c["sample"], "value_1_thread": c1["value"],
                                "note": "CPU port of the frozen spec (oracle/); the reference has no "
                                        "code for this path"}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bytes", type=int, default=16 * GIB, help="bytes per array (default 16 GiB)")
    ap.add_argument("--cpu-bytes", type=int, default=256 << 20, help="bytes per array of the CPU sample")
    ap.add_argument("--variant", default="auto", choices=["auto", "vec", "tma"])
    ap.add_argument("--e2e-calls", type=int, default=3)
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the e2e leg")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-fabric", action="store_true", help="skip the P2P / all-reduce legs at --gpus >= 2")
    ap.add_argument("--fabric-bytes", type=int, default=GIB, help="bytes per transfer of the headline P2P matrices")
    args = ap.parse_args()
    if args.steps < 1:
        raise SystemExit("--steps must be >= 1")
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
