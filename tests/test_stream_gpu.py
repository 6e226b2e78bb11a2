"""GPU parity: HBM copy / triad through the C ABI vs the CPU oracle (bit-exact fp32:
triad is one fused multiply-add per element on both sides), both kernel variants,
ragged sizes, and the whole-device probe verdict."""
import numpy as np
import pytest

import gpuprobe
import oracle_lib

pytestmark = pytest.mark.gpu
DEV = "0"
COPY, TRIAD = gpuprobe.GP_PROBE_COPY, gpuprobe.GP_PROBE_TRIAD

SIZES = [32, 64, 32 * 1023, 16 * 1024, 32 * 1024, 32 * 1024 * 7 + 96, (1 << 20) * 5 + 32 * 17,
         (1 << 20) * 64]


@pytest.fixture(params=["tma", "vec"])
def P(request, probe, probe_vec):
    return probe if request.param == "tma" else probe_vec


@pytest.mark.parametrize("nbytes", SIZES)
def test_copy_is_bit_exact(P, nbytes):
    rng = np.random.default_rng(nbytes)
    src = rng.integers(0, 256, size=nbytes, dtype=np.uint8)
    dst = np.zeros(nbytes, dtype=np.uint8)
    P.stream_host(DEV, COPY, src, None, 0.0, dst)
    assert np.array_equal(src, dst)


@pytest.mark.parametrize("nbytes", SIZES)
def test_triad_matches_oracle_bitwise(P, oracle, nbytes):
    n = nbytes // 4
    b = oracle_lib.gen_f32(oracle, 11, 0, n)
    c = oracle_lib.gen_f32(oracle, 12, 0, n)
    # stress rounding: scale some elements so that mul+add and fma differ
    b[::7] *= np.float32(1e-3)
    c[::5] *= np.float32(1.000123)
    a = np.zeros(n, dtype=np.float32)
    ref = np.zeros(n, dtype=np.float32)
    s = 3.0
    P.stream_host(DEV, TRIAD, b, c, s, a)
    oracle.gp_cpu_triad(ref.ctypes.data, b.ctypes.data, c.ctypes.data, s, n)
    assert np.array_equal(a.view(np.uint32), ref.view(np.uint32))


def test_triad_special_values(P, oracle):
    n = 8 * 1024
    b = np.zeros(n, dtype=np.float32)
    c = np.zeros(n, dtype=np.float32)
    b[:8] = [np.inf, -np.inf, 0.0, -0.0, 1e-45, 3.4e38, 1.0, -1.0]
    c[:8] = [1.0, 1.0, -0.0, 0.0, 1e-45, 3.4e38, 1e-45, 1.0 / 3.0]
    a = np.zeros(n, dtype=np.float32)
    ref = np.zeros(n, dtype=np.float32)
    P.stream_host(DEV, TRIAD, b, c, 3.0, a)
    oracle.gp_cpu_triad(ref.ctypes.data, b.ctypes.data, c.ctypes.data, 3.0, n)
    assert np.array_equal(a.view(np.uint32), ref.view(np.uint32))  # denormals kept, no flush-to-zero


def test_stream_probe_self_checks(P):
    nbytes = 256 << 20
    t, chk = P.stream_probe(DEV, COPY, nbytes, warmup=1, iters=3)
    n = nbytes // 8
    assert chk == (4 * n * (n - 1)) & ((1 << 64) - 1)  # closed form of the address pattern's sum
    assert t.launches == 3 and t.best_ms > 0
    t, bad = P.stream_probe(DEV, TRIAD, nbytes, warmup=1, iters=3)
    assert bad == 0 and t.best_ms > 0


def test_triad_generated_operands_match_oracle(probe, oracle):
    """gp_hbm_step's triad output checksum equals the oracle's for the same generator."""
    nbytes = 8 << 20
    out = probe.hbm_step(DEV, nbytes, step=5)
    n = nbytes // 4
    ref = np.empty(n, dtype=np.float32)
    oracle.gp_cpu_triad_gen(ref.ctypes.data, 0x5EED, 3.0, 0, n)
    chk = oracle_lib.verify(oracle, ref.view(np.uint64), 0, 0)  # sum/xor/wsum are kind-independent
    assert (out[2].sum, out[2].xor_, out[2].wsum) == (chk.sum, chk.xor_, chk.wsum)
    assert out[2].mismatches == 0
    nw = nbytes // 8
    r0, _ = oracle_lib.pattern(oracle, nw, 0, 5)
    r1, _ = oracle_lib.pattern(oracle, nw, 2, 1)
    assert out[0].astuple() == r0.astuple() and out[1].astuple() == r1.astuple()
    again = probe.hbm_step(DEV, nbytes, step=5)
    assert [r.astuple() for r in again] == [r.astuple() for r in out]  # idempotent
    stats = probe.kernel_stats(DEV)
    assert stats[gpuprobe.GP_K_VERIFY].launches >= 6 and stats[gpuprobe.GP_K_COPY].launches >= 2


def test_probe_device_verdict(oracle):
    """The call the plugin makes (device/devices.go:74 replacement), small sizes."""
    import ctypes as C
    with gpuprobe.Probe(hbm_bytes=256 << 20, pattern_bytes=64 << 20, wo_passes=8, iters=3, warmup=1) as p:
        mask = gpuprobe.GP_PROBE_COPY | gpuprobe.GP_PROBE_TRIAD | gpuprobe.GP_PROBE_PATTERN
        r = p.probe_device(p.device_uuid(0), mask)
        assert r.healthy == 1 and r.failed_mask == 0 and r.probed_mask == mask
        assert r.copy_ok == 1 and r.triad_ok == 1 and r.pattern_passes == 10 and r.pattern_mismatches == 0
        assert r.copy_gbps_best > 100 and r.triad_gbps_best > 100 and r.verify_gbps_best > 100
        assert r.cc_major == 10 and r.sm_count > 100
        nw = (64 << 20) // 8
        d = 0
        for j, (kind, pas) in enumerate([(0, q) for q in range(8)] + [(2, 0), (2, 1)]):
            rr, _ = oracle_lib.pattern(oracle, nw, kind, pas, nthreads=4)
            d = oracle.gp_cpu_digest_step(d, C.byref(rr), j)
        assert r.pattern_digest == d
        # an impossible floor turns the verdict Unhealthy without an error
    with gpuprobe.Probe(hbm_bytes=64 << 20, iters=2, warmup=1, min_copy_gbps=1e9) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_COPY)
        assert r.healthy == 0 and r.failed_mask == gpuprobe.GP_PROBE_COPY and r.error == 0


def test_triad_golden_fixture_on_gpu(probe):
    import json, os
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "stream_gemm_golden.json")))
    for t in gold["triad"]:
        out = probe.hbm_step(DEV, t["nbytes"], step=0)
        assert (hex(out[2].sum), hex(out[2].xor_), hex(out[2].wsum)) == (t["sum"], t["xor"], t["wsum"])


def test_triad_known_answer_from_the_cpu_oracle(oracle):
    """triad_ok is the GPU agreeing with itself (same fmaf, same chip): fine for memory faults, blind to a
    broken FMA.  The known answer closes that: SUM / XOR of the result's words computed OFF the device by the
    oracle's streaming gp_cpu_triad_checksum, handed in through gp_config, compared by the probe."""
    nbytes = 64 << 20
    esum, exor = oracle_lib.triad_checksum(oracle, 0x5EED, nbytes // 4, 3.0, nthreads=4)
    kw = dict(hbm_bytes=nbytes, iters=2, warmup=1)
    with gpuprobe.Probe(**kw) as p:                                     # nothing configured: reported, not compared
        r = p.probe_device("0", gpuprobe.GP_PROBE_TRIAD)
        assert (r.triad_sum, r.triad_xor) == (esum, exor) and r.triad_known_answer == 0 and r.healthy == 1
    with gpuprobe.Probe(triad_expect_sum=esum, triad_expect_xor=exor, **kw) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_TRIAD)
        assert r.triad_known_answer == 1 and r.triad_ok == 1 and r.healthy == 1
    with gpuprobe.Probe(triad_expect_sum=esum ^ 1, triad_expect_xor=exor, **kw) as p:   # a device whose FMA is off by a bit
        r = p.probe_device("0", gpuprobe.GP_PROBE_TRIAD)
        assert r.triad_known_answer == 2 and r.triad_ok == 1 and r.healthy == 0
        assert r.failed_mask == gpuprobe.GP_PROBE_TRIAD


def test_triad_known_answer_default_configuration():
    """The 16 GiB / seed 0x5EED constant committed in tests/golden/triad_known_answers.json (made by
    tests/golden/make_triad_known_answers.py on the CPU) against the device at the full default size."""
    import json
    import os
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "triad_known_answers.json")))
    k = gold["answers"]["17179869184"]
    with gpuprobe.Probe(iters=2, warmup=1, triad_expect_sum=int(k["sum"], 16), triad_expect_xor=int(k["xor"], 16)) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_TRIAD)
        assert r.hbm_bytes == 16 << 30 and r.triad_known_answer == 1 and r.healthy == 1
