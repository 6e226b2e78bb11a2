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
        with pytest.raises(gpuprobe.GpuProbeError):
            p.nvml_snapshot("GPU-00000000-0000-0000-0000-000000000000")


def test_whole_memory_grading():
    """pattern_bytes = GP_PATTERN_ALL_FREE: fill/verify every free byte but a 1 GiB margin."""
    with gpuprobe.Probe(pattern_bytes=gpuprobe.GP_PATTERN_ALL_FREE, wo_passes=2) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_PATTERN)
        assert r.healthy == 1 and r.pattern_passes == 4 and r.pattern_mismatches == 0
        assert r.pattern_bytes > (100 << 30) and r.pattern_bytes % (1 << 20) == 0  # most of the 180 GB card
        assert r.verify_gbps_best > 1000 and r.fill_gbps_best > 1000
