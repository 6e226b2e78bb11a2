"""GPU parity: pattern fill/verify through the C ABI vs the CPU oracle (bit-exact),
golden fixtures, closed forms at BASELINE.json's full size, fault injection."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import gpuprobe
import oracle_lib
from test_oracle import closed_sum_xor, M64

pytestmark = pytest.mark.gpu
WO, WZ, AD = 0, 1, 2
DEV = "0"


def as_tuple(r):
    return r.astuple()


@pytest.mark.parametrize("nwords", [64, 128, 4096, 64 * 1021, 1 << 20, (1 << 22) + 64 * 3])
@pytest.mark.parametrize("kind,pas", [(WO, 0), (WO, 5), (WO, 63), (WZ, 31), (AD, 0), (AD, 1)])
def test_fill_verify_matches_oracle(probe, oracle, nwords, kind, pas):
    host = np.zeros(nwords, dtype=np.uint64)
    r = probe.pattern_host(DEV, host, kind, pas, fill=True)  # device fills, verifies, copies back
    ref, buf = oracle_lib.pattern(oracle, nwords, kind, pas, nthreads=4)
    assert as_tuple(r) == as_tuple(ref)
    assert np.array_equal(host, buf)  # every word the fill kernel wrote is the oracle's word


def test_empty_input(probe):
    r = probe.pattern_host(DEV, None, WO, 0, fill=True)
    assert as_tuple(r) == (0, 0, 0, 0, M64, 0)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_verify_arbitrary_host_data(probe, oracle, seed):
    """Verify of words the device did not write: random data, every word a mismatch."""
    rng = np.random.default_rng(seed)
    n = 64 * 4099
    w = rng.integers(0, 1 << 63, size=n, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=n, dtype=np.uint64)
    for kind, pas in [(WO, 7), (AD, 1)]:
        r = probe.pattern_host(DEV, w, kind, pas, fill=False)
        ref = oracle_lib.verify(oracle, w, kind, pas, nthreads=4)
        assert as_tuple(r) == as_tuple(ref)
        assert r.mismatches > n - 8


@pytest.mark.parametrize("kind,pas", [(WO, 3), (AD, 0), (AD, 1), (WZ, 9)])
def test_bitflip_fault_injection(probe, oracle, kind, pas):
    n = 1 << 20
    probe.pattern_fill(DEV, kind, pas, n)
    clean = probe.pattern_verify(DEV, kind, pas, n)
    assert clean.mismatches == 0
    for j, b in [(0, 0), (n - 1, 63), (123457, 17)]:
        probe.inject_bitflip(DEV, j, b)
        r = probe.pattern_verify(DEV, kind, pas, n)
        assert (r.mismatches, r.first_bad_index, r.first_bad_xor) == (1, j, 1 << b)
        assert r.xor_ == clean.xor_ ^ (1 << b)
        assert (r.sum - clean.sum) & M64 in ((1 << b), (-(1 << b)) & M64)
        probe.inject_bitflip(DEV, j, b)  # flip back
    # two faults: the smaller index is reported, whichever CTA saw it
    probe.inject_bitflip(DEV, 900001, 1)
    probe.inject_bitflip(DEV, 300, 2)
    r = probe.pattern_verify(DEV, kind, pas, n)
    assert (r.mismatches, r.first_bad_index, r.first_bad_xor) == (2, 300, 4)
    # and the same two faults through the oracle
    _, buf = oracle_lib.pattern(oracle, n, kind, pas, nthreads=4)
    buf[900001] ^= np.uint64(2)
    buf[300] ^= np.uint64(4)
    assert as_tuple(r) == as_tuple(oracle_lib.verify(oracle, buf, kind, pas, nthreads=4))
    assert probe.pattern_verify(DEV, kind, pas, n).astuple() == as_tuple(r)  # idempotent


def test_random_single_faults_are_always_found_exactly(probe):
    """Property: one flipped bit anywhere in 1 GiB, any pattern, any pass, is reported with its
    exact word index and xor, and sum/xor move by exactly that bit."""
    rng = np.random.default_rng(2026)
    n = (1 << 30) // 8
    for _ in range(24):
        kind = int(rng.choice([WO, WZ, AD]))
        pas = int(rng.integers(0, 2 if kind == AD else 64))
        j = int(rng.integers(0, n))
        b = int(rng.integers(0, 64))
        probe.pattern_fill(DEV, kind, pas, n)
        s, x = closed_sum_xor(kind, pas, n)
        probe.inject_bitflip(DEV, j, b)
        r = probe.pattern_verify(DEV, kind, pas, n)
        assert (r.mismatches, r.first_bad_index, r.first_bad_xor) == (1, j, 1 << b), (kind, pas, j, b)
        assert r.xor_ == x ^ (1 << b) and (r.sum - s) & M64 in ((1 << b), (-(1 << b)) & M64)


def test_wrong_pass_is_all_mismatch(probe, oracle):
    n = 1 << 16
    probe.pattern_fill(DEV, WO, 0, n)
    r = probe.pattern_verify(DEV, WO, 1, n)
    assert r.mismatches == n and r.first_bad_index == 0 and r.first_bad_xor == 3


def test_golden_fixtures_on_gpu(probe):
    path = os.path.join(os.path.dirname(__file__), "golden", "pattern_golden.json")
    gold = json.load(open(path))
    for case in gold["pattern"]:
        r = probe.pattern_host(DEV, None, case["kind"], case["pass"], fill=True) if case["nwords"] == 0 else None
        probe.pattern_fill(DEV, case["kind"], case["pass"], case["nwords"])
        r = probe.pattern_verify(DEV, case["kind"], case["pass"], case["nwords"])
        assert [hex(v) for v in as_tuple(r)] == case["result"], case


def test_pattern_probe_all_passes_and_digest(probe, oracle):
    """64 walking-ones passes + 2 address passes in one call; checksum of checksums vs golden."""
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "pattern_golden.json")))
    case = [c for c in gold["digest"] if c["nwords"] == 1 << 18][0]
    res, ft, vt = probe.pattern_probe(DEV, WO, 0, 64, 1 << 18)
    res2, _, _ = probe.pattern_probe(DEV, AD, 0, 2, 1 << 18)
    d = 0
    for j, r in enumerate(res + res2):
        rr = oracle_lib.PatternResult(*as_tuple(r))
        d = oracle.gp_cpu_digest_step(d, C.byref(rr), j)
        assert r.mismatches == 0
    assert hex(d) == case["digest"]
    assert ft.launches == 64 and vt.launches == 64 and vt.best_ms > 0


def test_full_size_closed_forms(probe):
    """BASELINE.json configs[1]: 16 GiB = 2^31 words.  Size-independent properties: closed-form
    sum/xor, zero mismatches, and a single injected fault found exactly."""
    n = 1 << 31
    for kind, pas in [(WO, 0), (WO, 41), (AD, 0), (AD, 1)]:
        probe.pattern_fill(DEV, kind, pas, n)
        r = probe.pattern_verify(DEV, kind, pas, n)
        s, x = closed_sum_xor(kind, pas, n)
        assert (r.sum, r.xor_, r.mismatches, r.first_bad_index) == (s, x, 0, M64), (kind, pas)
    assert closed_sum_xor(WO, 0, n)[0] == 0xFFFFFFFFFE000000
    assert closed_sum_xor(AD, 0, n)[0] == 0xFFFFFFFE00000000
    j = (1 << 31) - 12345
    probe.inject_bitflip(DEV, j, 62)
    r = probe.pattern_verify(DEV, AD, 1, n)
    assert (r.mismatches, r.first_bad_index, r.first_bad_xor) == (1, j, 1 << 62)
    probe.trim()


def test_maximum_size_96GiB_indices_past_2_pow_32(probe):
    """Half the card in one launch: 96 GiB = 1.29e10 words, word indices need 34 bits.  Closed forms
    (mod 2^64) for sum and xor, no mismatches, and a fault planted above index 2^33 found exactly."""
    probe.trim()
    with gpuprobe.Probe(keep_arena=0, arena_cap_bytes=100 << 30) as big:
        n = (96 << 30) // 8
        for kind, pas in [(WO, 63), (AD, 0)]:
            big.pattern_fill(DEV, kind, pas, n)
            r = big.pattern_verify(DEV, kind, pas, n)
            s, x = closed_sum_xor(kind, pas, n)
            assert (r.sum, r.xor_, r.mismatches, r.first_bad_index) == (s, x, 0, M64), (kind, pas)
        j = (1 << 33) + 987654321
        big.inject_bitflip(DEV, j, 5)
        big.inject_bitflip(DEV, n - 1, 63)
        r = big.pattern_verify(DEV, AD, 0, n)
        assert (r.mismatches, r.first_bad_index, r.first_bad_xor) == (2, j, 1 << 5)


def test_invalid_arguments(probe):
    for args in [(WO, 64, 64), (AD, 2, 64), (7, 0, 64), (WO, 0, 65)]:
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            probe.pattern_fill(DEV, *args)
        assert e.value.code == gpuprobe.GP_ERR_INVALID
    with pytest.raises(gpuprobe.GpuProbeError) as e:
        probe.pattern_fill("GPU-00000000-0000-0000-0000-000000000000", WO, 0, 64)
    assert e.value.code == gpuprobe.GP_ERR_NO_DEVICE


def test_uuid_addressing(probe):
    u = probe.device_uuid(0)
    assert u.startswith("GPU-") and len(u) == 40
    probe.pattern_fill(u, WO, 0, 4096)
    assert probe.pattern_verify(u.upper().replace("GPU-", "GPU-"), WO, 0, 4096).mismatches == 0
