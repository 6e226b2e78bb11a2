"""Engine behaviour through the C ABI: locking, arena limits, lifecycle, probe_all."""
import threading

import numpy as np
import pytest

import gpuprobe

pytestmark = pytest.mark.gpu


def test_same_device_calls_serialise_and_stay_correct():
    """cgo calls land on arbitrary OS threads: hammer one device from four threads."""
    with gpuprobe.Probe(keep_arena=1) as p:
        errs = []

        def work(seed):
            try:
                for k in range(6):
                    n = 64 * (500 + 37 * seed + k)
                    host = np.zeros(n, dtype=np.uint64)
                    r = p.pattern_host("0", host, gpuprobe.GP_PAT_ADDRESS, seed & 1, fill=True)
                    exp = np.arange(n, dtype=np.uint64) * np.uint64(8)
                    if seed & 1:
                        exp = ~exp
                    assert r.mismatches == 0 and np.array_equal(host, exp)
            except Exception as e:  # noqa: BLE001
                errs.append(e)

        ts = [threading.Thread(target=work, args=(s,)) for s in range(4)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert not errs, errs


def test_arena_cap_is_enforced_and_reported():
    with gpuprobe.Probe(arena_cap_bytes=64 << 20) as p:
        p.pattern_fill("0", 0, 0, (32 << 20) // 8)
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            p.pattern_fill("0", 0, 0, (128 << 20) // 8)
        assert e.value.code == gpuprobe.GP_ERR_NOMEM and "exceeds cap" in e.value.detail
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            p.stream_probe("0", gpuprobe.GP_PROBE_TRIAD, 64 << 20, 1, 1)
        assert e.value.code == gpuprobe.GP_ERR_NOMEM
        # a device result carries the error and is not healthy
        with pytest.raises(gpuprobe.GpuProbeError):
            p.probe_device("0", gpuprobe.GP_PROBE_COPY)


def test_trim_releases_and_reallocates():
    import ctypes as C
    with gpuprobe.Probe(keep_arena=1) as p:
        p.pattern_fill("0", 0, 0, (256 << 20) // 8)
        p.trim()
        with pytest.raises(gpuprobe.GpuProbeError) as e:  # verify after trim: nothing to verify
            p.pattern_verify("0", 0, 0, 64)
        assert e.value.code == gpuprobe.GP_ERR_INVALID
        p.pattern_fill("0", 0, 0, 64)
        assert p.pattern_verify("0", 0, 0, 64).mismatches == 0


def test_probe_all_one_gpu_and_bad_counts():
    with gpuprobe.Probe(hbm_bytes=64 << 20, pattern_bytes=16 << 20, wo_passes=2, iters=2, warmup=1, gemm_mnk=512) as p:
        res, wall = p.probe_all(1)
        assert len(res) == 1 and res[0].healthy == 1 and wall > 0
        assert res[0].probed_mask == gpuprobe.GP_PROBE_ALL_SINGLE and res[0].gemm.samples_bad == 0
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            p.probe_all(p.device_count() + 1)
        assert e.value.code == gpuprobe.GP_ERR_NO_DEVICE
        ident = p.probe_device("0", 0)  # mask 0: identity only, nothing launched
        assert ident.probed_mask == 0 and ident.healthy == 1 and ident.uuid.startswith(b"GPU-")


def test_two_contexts_coexist():
    with gpuprobe.Probe(keep_arena=1) as a, gpuprobe.Probe(keep_arena=1) as b:
        a.pattern_fill("0", 0, 3, 4096)
        b.pattern_fill("0", 2, 1, 4096)
        assert a.pattern_verify("0", 0, 3, 4096).mismatches == 0
        assert b.pattern_verify("0", 2, 1, 4096).mismatches == 0


def test_nvml_snapshot():
    with gpuprobe.Probe() as p:
        i = p.nvml_snapshot(p.device_uuid(0))
        if not i.available:
            pytest.skip("NVML not loadable on this box")
        assert 100 < i.sm_clock_max_mhz < 4000 and i.sm_clock_mhz <= i.sm_clock_max_mhz
        assert i.mem_total_bytes > (100 << 30) and i.power_mw > 1000 and 0 < i.temperature_c < 110
        assert i.other_compute_procs == 0  # a fresh box: nobody else is computing here
        # ... and still nobody once THIS process holds a context on the device: NVML lists host pids, which
        # getpid() does not match inside a container, and the library must not count itself as a tenant
        p.pattern_fill(p.device_uuid(0), 0, 0, 64)
        assert p.nvml_snapshot(p.device_uuid(0)).other_compute_procs == 0
        with pytest.raises(gpuprobe.GpuProbeError):
            p.nvml_snapshot("GPU-00000000-0000-0000-0000-000000000000")


def test_whole_memory_grading():
    """pattern_bytes = GP_PATTERN_ALL_FREE: fill/verify every free byte but a 1 GiB margin."""
    with gpuprobe.Probe(pattern_bytes=gpuprobe.GP_PATTERN_ALL_FREE, wo_passes=2) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_PATTERN)
        assert r.healthy == 1 and r.pattern_passes == 4 and r.pattern_mismatches == 0
        assert r.pattern_bytes > (100 << 30) and r.pattern_bytes % (1 << 20) == 0  # most of the 180 GB card
        assert r.verify_gbps_best > 1000 and r.fill_gbps_best > 1000
    # with every probe selected the GEMM shares the one big arena (no second buffer beside a card-sized one)
    with gpuprobe.Probe(pattern_bytes=gpuprobe.GP_PATTERN_ALL_FREE, wo_passes=1, hbm_bytes=1 << 30, iters=2, warmup=1,
                        gemm_mnk=2048) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_ALL_SINGLE)
        assert r.healthy == 1 and r.probed_mask == gpuprobe.GP_PROBE_ALL_SINGLE and r.pattern_bytes > (100 << 30)


@pytest.mark.parametrize("backend", [gpuprobe.GP_ARENA_VMM, gpuprobe.GP_ARENA_MALLOC, gpuprobe.GP_ARENA_AUTO],
                         ids=["vmm", "cudaMalloc", "auto"])
def test_arena_time_is_reported_and_a_kept_arena_costs_nothing(backend):
    """gp_device_result says where the call's time went.  keep_arena = 0: every call grows the arena
    (alloc_s > 0) and hands it to the reaper thread on return (free_deferred, free_s = 0); the next call
    waits for that release before it allocates.  keep_arena = 1 (the periodic re-probe of an idle device): the
    second call finds the arena in place — no allocation inside the call — until gp_trim_device hands it back."""
    mask = gpuprobe.GP_PROBE_COPY | gpuprobe.GP_PROBE_PATTERN
    kw = dict(hbm_bytes=1 << 30, wo_passes=2, iters=2, warmup=1, arena_backend=backend)
    used = {gpuprobe.GP_ARENA_AUTO: gpuprobe.GP_ARENA_MALLOC}.get(backend, backend)  # AUTO, cold every call: cudaMalloc
    with gpuprobe.Probe(keep_arena=0, **kw) as p:
        for _ in range(2):
            r = p.probe_device("0", mask)
            assert r.healthy == 1 and r.arena_backend == used and r.arena_reused == 0
            assert r.alloc_s > 0 and r.free_s == 0 and r.free_deferred == 1 and r.arena_bytes >= 2 << 30
            assert abs(r.duration_s - (r.alloc_s + r.free_s + r.kernels_s)) < 1e-9
        # all four probes cold: the GEMM runs in its own small buffer while the arena is allocated behind it
        r = p.probe_device("0", gpuprobe.GP_PROBE_ALL_SINGLE)
        assert r.healthy == 1 and r.probed_mask == gpuprobe.GP_PROBE_ALL_SINGLE and r.gemm.samples_bad == 0
        p.trim()   # joins the reaper: the memory is back
        import torch
        free, total = torch.cuda.mem_get_info(0)
        assert free > total - (24 << 30)   # nothing of ours is left (the session fixtures keep a few GiB)
    used = {gpuprobe.GP_ARENA_AUTO: gpuprobe.GP_ARENA_VMM}.get(backend, backend)     # AUTO, kept arena: VMM
    with gpuprobe.Probe(keep_arena=1, **kw) as p:
        first = p.probe_device("0", mask)
        again = p.probe_device("0", mask)
        assert first.alloc_s > 0 and first.arena_reused == 0 and first.free_s == 0 and first.arena_backend == used
        assert again.free_deferred == 0
        assert again.alloc_s == 0 and again.free_s == 0 and again.arena_reused == 1 and again.healthy == 1
        # growing a kept arena maps only the difference and leaves the lower part where it was
        p.pattern_fill("0", gpuprobe.GP_PAT_ADDRESS, 0, (1 << 30) // 8)
        p.pattern_fill("0", gpuprobe.GP_PAT_WALKING_ONES, 5, (3 << 30) // 8)
        assert p.pattern_verify("0", gpuprobe.GP_PAT_WALKING_ONES, 5, (3 << 30) // 8).mismatches == 0
        p.trim_device("0")   # what the plugin calls from Allocate(): the pod gets the memory back
        after = p.probe_device("0", mask)
        assert after.arena_reused == 0 and after.alloc_s > 0 and after.healthy == 1


def test_sm_limit_runs_the_launchers_at_slice_scale():
    """A 1g MIG slice exposes ~1/7 of the SMs and 23 GiB: the grids and the arena must size themselves to
    what the device reports.  gp_config.sm_limit makes the launchers believe in 18 SMs (9 CTA pairs)."""
    with gpuprobe.Probe(hbm_bytes=256 << 20, wo_passes=2, iters=2, warmup=1, gemm_mnk=2048, sm_limit=18) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_ALL_SINGLE)
        assert r.healthy == 1 and r.sm_count == 18 and r.gemm.samples_bad == 0
        full = gpuprobe.Probe(hbm_bytes=256 << 20, gemm_mnk=2048, iters=2, warmup=1)
        g_full = full.gemm_probe("0", mnk=2048, iters=3, warmup=1, samples=256)
        full.close()
        g = p.gemm_probe("0", mnk=2048, iters=3, warmup=1, samples=256)
        assert g.samples_bad == 0 and g.tflops_best < 0.5 * g_full.tflops_best  # really ran on fewer CTAs
    with gpuprobe.Probe(sm_limit=1, gemm_mnk=512, gemm_variant=gpuprobe.GP_GEMM_AUTO) as p:  # odd / tiny counts
        assert p.gemm_probe("0", mnk=512, iters=1, warmup=0, samples=64).samples_bad == 0


_POISON_SCRIPT = r"""
import sys
sys.path.insert(0, sys.argv[1])
import gpuprobe
p = gpuprobe.Probe(hbm_bytes=64 << 20, wo_passes=1, iters=1, warmup=1)
dev = p.device_uuid(0)
mask = gpuprobe.GP_PROBE_COPY | gpuprobe.GP_PROBE_PATTERN
assert p.probe_device(dev, mask).healthy == 1
assert not p.device_poisoned(dev)
# a non-sticky error (2 = cudaErrorMemoryAllocation) is reported and forgotten
assert p.inject_cuda_error(dev, 2) == gpuprobe.GP_ERR_NOMEM and not p.device_poisoned(dev)
# a sticky one (700 = cudaErrorIllegalAddress) loses the device for this process
assert p.inject_cuda_error(dev, 700) == gpuprobe.GP_ERR_DEVICE_LOST and p.device_poisoned(dev)
for call in (lambda: p.probe_device(dev, mask), lambda: p.pattern_fill(dev, 0, 0, 64)):
    try:
        call()
        raise SystemExit("a call on a lost device succeeded")
    except gpuprobe.GpuProbeError as e:
        assert e.code == gpuprobe.GP_ERR_DEVICE_LOST, e
        assert "gp_device_reset" in e.detail
p.device_reset(dev)            # in-process recovery: only for hosts where nothing else uses CUDA
assert not p.device_poisoned(dev)
r = p.probe_device(dev, mask)
assert r.healthy == 1 and r.alloc_s > 0
p.close()
print("poison-reset ok")
"""


def test_sticky_fault_marks_the_device_lost_and_reset_recovers(tmp_path):
    """Runs in a child process (gp_device_reset is cudaDeviceReset: it would take this pytest process's
    torch context with it).  The fault itself is injected at the API level — nothing abnormal reaches the GPU."""
    import os
    import subprocess
    import sys
    pkg = os.path.dirname(os.path.dirname(gpuprobe.__file__))
    r = subprocess.run([sys.executable, "-c", _POISON_SCRIPT, pkg], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "poison-reset ok" in r.stdout, (r.stdout, r.stderr)


def test_cold_probe_that_runs_out_of_memory_cleans_up():
    """Three 80 GiB slabs do not fit a 180 GB card: the third allocation fails on the provider thread while the
    call is already working.  The call must come back with GP_ERR_NOMEM (not hang, not crash), and the two slabs
    it did get must be back with the driver afterwards — the next, ordinary probe succeeds."""
    import torch
    with gpuprobe.Probe(hbm_bytes=80 << 30, iters=1, warmup=1, arena_cap_bytes=400 << 30) as p:
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            p.probe_device("0", gpuprobe.GP_PROBE_TRIAD)
        assert e.value.code == gpuprobe.GP_ERR_NOMEM
        p.trim()  # joins the release thread
        free, total = torch.cuda.mem_get_info(0)
        assert free > total - (24 << 30)
    with gpuprobe.Probe(hbm_bytes=1 << 30, iters=1, warmup=1, wo_passes=1) as p:
        assert p.probe_device("0", gpuprobe.GP_PROBE_ALL_SINGLE).healthy == 1
