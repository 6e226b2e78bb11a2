"""Pins the CPU oracle (oracle/gpuprobe_oracle.c) to the closed-form known answers of
SURVEY.md §8(c), to a brute-force pure-Python restatement on small N, and to the
committed golden fixtures.  The reference has no tests or vectors for this path
("parity unpinned"), so these are the only anchors there are.
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_lib

M64 = (1 << 64) - 1
WO, WZ, AD = 0, 1, 2


def py_word(kind, pas, i):
    if kind == WO:
        return 1 << ((i + pas) & 63)
    if kind == WZ:
        return ~(1 << ((i + pas) & 63)) & M64
    a = (8 * i) & M64
    return (~a & M64) if (pas & 1) else a


def py_verify(words, kind, pas):
    s = x = ws = mism = 0
    first, fx = M64, 0
    for i, w in enumerate(words):
        w = int(w)
        s = (s + w) & M64
        x ^= w
        ws = (ws + w * (2 * i + 1)) & M64
        e = py_word(kind, pas, i)
        if w != e:
            if mism == 0:
                first, fx = i, w ^ e
            mism += 1
    return (s, x, ws, mism, first, fx)


# ---- closed forms (SURVEY.md §8c "Known answers") -----------------------------------------
def closed_sum_xor(kind, pas, n):
    if kind == WO:
        return (-(n // 64)) & M64, 0 if (n // 64) % 2 == 0 else M64
    if kind == WZ:  # complement of walking-ones: sum = N*(2^64-1) - sum_wo, xor flips when N odd
        s_wo = (-(n // 64)) & M64
        return (n * M64 - s_wo) & M64, 0 if (n // 64) % 2 == 0 else M64
    big_x = lambda m: [m, 1, m + 1, 0][m % 4]
    s, x = (4 * n * (n - 1)) & M64, (8 * big_x(n - 1)) & M64
    if pas & 1:
        # ~a = (2^64-1) - a ; N complemented words
        s = (n * M64 - s) & M64
        x = x ^ (M64 if n % 2 else 0)
    return s, x


@pytest.mark.parametrize("nwords,kind,pas,sum_,xor_", [
    (1 << 25, WO, 0, 0xFFFFFFFFFFF80000, 0),
    (1 << 25, WO, 37, 0xFFFFFFFFFFF80000, 0),
    (1 << 25, AD, 0, 0x000FFFFFF8000000, 0),
])
def test_known_answers_256MiB(oracle, nwords, kind, pas, sum_, xor_):
    r, _ = oracle_lib.pattern(oracle, nwords, kind, pas, nthreads=8)
    assert r.sum == sum_ and r.xor_ == xor_
    assert r.mismatches == 0 and r.first_bad_index == M64 and r.first_bad_xor == 0


def test_known_answers_16GiB_closed_form_only():
    # 2^31 words is not run on the CPU here; the closed forms must give SURVEY's table.
    assert closed_sum_xor(WO, 0, 1 << 31) == (0xFFFFFFFFFE000000, 0)
    assert closed_sum_xor(AD, 0, 1 << 31) == (0xFFFFFFFE00000000, 0)


@pytest.mark.parametrize("pas,wsum", [(0, 0xFFFFFFFFFFFC10C0), (5, 0xFFFFFFFFFFFFF340)])
def test_known_wsum_4096(oracle, pas, wsum):
    r, _ = oracle_lib.pattern(oracle, 4096, WO, pas)
    assert r.wsum == wsum


@pytest.mark.parametrize("kind,passes", [(WO, [0, 1, 5, 63]), (WZ, [0, 17, 63]), (AD, [0, 1])])
@pytest.mark.parametrize("nwords", [0, 64, 128, 4096, 64 * 37])
def test_oracle_vs_python_bruteforce(oracle, kind, passes, nwords):
    for pas in passes:
        for nt in (1, 3):
            r, buf = oracle_lib.pattern(oracle, nwords, kind, pas, nthreads=nt)
            assert r.astuple() == py_verify(buf, kind, pas)
            assert [int(w) for w in buf[:130]] == [py_word(kind, pas, i) for i in range(min(nwords, 130))]
            if nwords:
                s, x = closed_sum_xor(kind, pas, nwords)
                assert (r.sum, r.xor_) == (s, x)


@pytest.mark.parametrize("kind,pas", [(WO, 3), (AD, 0), (AD, 1), (WZ, 9)])
def test_bitflip_negative(oracle, kind, pas):
    """SURVEY.md §8c negative test: flip bit b of word j."""
    n = 4096
    clean, buf = oracle_lib.pattern(oracle, n, kind, pas)
    for j, b in [(0, 0), (n - 1, 63), (1234, 17)]:
        w = buf.copy()
        w[j] ^= np.uint64(1 << b)
        r = oracle_lib.verify(oracle, w, kind, pas, nthreads=2)
        assert r.mismatches == 1 and r.first_bad_index == j and r.first_bad_xor == 1 << b
        assert r.xor_ == clean.xor_ ^ (1 << b)
        delta = (r.sum - clean.sum) & M64
        assert delta in ((1 << b), (-(1 << b)) & M64)
        assert r.astuple() == py_verify(w, kind, pas)
    # two flips: first index is the smaller one
    w = buf.copy()
    w[700] ^= np.uint64(2)
    w[300] ^= np.uint64(4)
    r = oracle_lib.verify(oracle, w, kind, pas, nthreads=4)
    assert (r.mismatches, r.first_bad_index, r.first_bad_xor) == (2, 300, 4)


def test_invalid_arguments(oracle):
    buf = np.zeros(64, dtype=np.uint64)
    r = oracle_lib.PatternResult()
    assert oracle.gp_cpu_pattern(buf.ctypes.data, 64, WO, 64, 1, C.byref(r)) != 0
    assert oracle.gp_cpu_pattern(buf.ctypes.data, 64, AD, 2, 1, C.byref(r)) != 0
    assert oracle.gp_cpu_pattern(buf.ctypes.data, 64, 7, 0, 1, C.byref(r)) != 0


# ---- generator / triad / GEMM restatements ---------------------------------------------------
def py_mix64(seed, idx):
    z = (seed + (idx + 1) * 0x9E3779B97F4A7C15) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def test_generator_matches_python(oracle):
    for seed in (0x5EED, 0, M64):
        for idx in (0, 1, 12345, (1 << 40) + 7):
            m = py_mix64(seed, idx)
            assert oracle.gp_cpu_mix64(seed, idx) == m
            f = ((m >> 40) - (1 << 23)) / float(1 << 23)
            assert oracle.gp_cpu_gen_f32(seed, idx) == np.float32(f)
            assert -1.0 <= f < 1.0
            # round-to-nearest-even bf16 of an exactly representable fp32
            bits = np.float32(f).view(np.uint32)
            rne = (int(bits) + 0x7FFF + ((int(bits) >> 16) & 1)) >> 16
            assert oracle.gp_cpu_gen_bf16(seed, idx) == rne & 0xFFFF


def test_triad_is_single_rounding(oracle):
    n = 4096
    b = oracle_lib.gen_f32(oracle, 1, 0, n)
    c = oracle_lib.gen_f32(oracle, 2, 0, n)
    a = np.empty(n, dtype=np.float32)
    oracle.gp_cpu_triad(a.ctypes.data, b.ctypes.data, c.ctypes.data, 3.0, n)
    exact = (b.astype(np.float64) + 3.0 * c.astype(np.float64)).astype(np.float32)  # one rounding
    assert np.array_equal(a.view(np.uint32), exact.view(np.uint32))
    g = np.empty(n, dtype=np.float32)
    oracle.gp_cpu_triad_gen(g.ctypes.data, 0x5EED, 3.0, 0, n)
    bb = oracle_lib.gen_f32(oracle, 0x5EED ^ oracle_lib.SEED_TRIAD_B_XOR, 0, n)
    cc = oracle_lib.gen_f32(oracle, 0x5EED ^ oracle_lib.SEED_TRIAD_C_XOR, 0, n)
    assert np.array_equal(g, (bb.astype(np.float64) + 3.0 * cc.astype(np.float64)).astype(np.float32))


def test_gemm_ref_matches_numpy(oracle):
    m, n, k = 16, 24, 64
    a = oracle_lib.gen_bf16(oracle, 0x5EED, 0, m * k)
    b = oracle_lib.gen_bf16(oracle, 0x5EED ^ oracle_lib.SEED_B_XOR, 0, n * k)
    c = np.empty((m, n), dtype=np.float64)
    ad = np.empty((m, n), dtype=np.float64)
    oracle.gp_cpu_gemm_ref(a.ctypes.data, b.ctypes.data, c.ctypes.data, ad.ctypes.data, m, n, k)
    af = (a.astype(np.uint32) << 16).view(np.float32).reshape(m, k).astype(np.float64)
    bf = (b.astype(np.uint32) << 16).view(np.float32).reshape(n, k).astype(np.float64)
    assert np.allclose(c, af @ bf.T, rtol=0, atol=1e-12)
    assert np.allclose(ad, np.abs(af) @ np.abs(bf).T, rtol=0, atol=1e-12)
    absdot = C.c_double()
    v = oracle.gp_cpu_gemm_sample(0x5EED, k, 3, 5, C.byref(absdot))
    assert abs(v - c[3, 5]) < 1e-12 and abs(absdot.value - ad[3, 5]) < 1e-12


# ---- golden fixtures (generated by tests/golden/make_golden.py from the oracle itself) -------
def test_golden_fixtures(oracle):
    path = os.path.join(os.path.dirname(__file__), "golden", "pattern_golden.json")
    with open(path) as f:
        gold = json.load(f)
    assert gold["source"].startswith("oracle")
    for case in gold["pattern"]:
        r, _ = oracle_lib.pattern(oracle, case["nwords"], case["kind"], case["pass"], nthreads=2)
        assert [hex(v) for v in r.astuple()] == case["result"], case
    for case in gold["digest"]:
        d = 0
        for j, (kind, pas) in enumerate(case["passes"]):
            r, _ = oracle_lib.pattern(oracle, case["nwords"], kind, pas)
            d = oracle.gp_cpu_digest_step(d, C.byref(r), j)
        assert hex(d) == case["digest"]
    g = gold["generator"]
    assert [int(x) for x in oracle_lib.gen_bf16(oracle, g["seed"], 0, len(g["bf16"]))] == g["bf16"]


def test_cpu_hbm_step_is_consistent(oracle):
    """The CPU port of the bench step agrees with the oracle's own building blocks."""
    nbytes = 1 << 20
    arena = np.zeros(5 * nbytes, dtype=np.uint8)
    out = oracle_lib.hbm_step(oracle, arena, nbytes, step=7, nthreads=3)
    nw = nbytes // 8
    r0, _ = oracle_lib.pattern(oracle, nw, WO, 7)
    r1, _ = oracle_lib.pattern(oracle, nw, AD, 1)
    assert out[0].astuple() == r0.astuple() and out[1].astuple() == r1.astuple()
    ref = np.empty(nbytes // 4, dtype=np.float32)
    oracle.gp_cpu_triad_gen(ref.ctypes.data, 0x5EED, 3.0, 0, nbytes // 4)
    chk = oracle_lib.verify(oracle, ref.view(np.uint64), WO, 0)
    assert (out[2].sum, out[2].xor_, out[2].wsum, out[2].mismatches) == (chk.sum, chk.xor_, chk.wsum, 0)
    z = arena[2 * nbytes:3 * nbytes].view(np.float32)
    assert np.array_equal(z, ref)


def test_golden_stream_gemm_fixtures(oracle):
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "stream_gemm_golden.json")))
    for t in gold["triad"]:
        n = t["nbytes"] // 4
        a = np.empty(n, dtype=np.float32)
        oracle.gp_cpu_triad_gen(a.ctypes.data, t["seed"], t["scalar"], 0, n)
        r = oracle_lib.verify(oracle, a.view(np.uint64), WO, 0)
        assert (hex(r.sum), hex(r.xor_), hex(r.wsum)) == (t["sum"], t["xor"], t["wsum"])
    for g in gold["gemm"]:
        for pt in g["points"]:
            ad = C.c_double()
            v = oracle.gp_cpu_gemm_sample(g["seed"], g["mnk"], pt["i"], pt["j"], C.byref(ad))
            assert v == pt["ref"] and ad.value == pt["absdot"]


def test_triad_known_answer_streaming_matches_stored_and_golden(oracle):
    """gp_cpu_triad_checksum (nothing stored) against the stored-array port of the step, and against the committed
    known answers (tests/golden/triad_known_answers.json) at the smallest size."""
    import json
    import os
    import oracle_lib
    n = 1 << 18
    arena = np.zeros(5 * n * 4, dtype=np.uint8)
    out = oracle_lib.hbm_step(oracle, arena, n * 4, step=0, nthreads=2)
    for threads in (1, 3, 8):
        assert oracle_lib.triad_checksum(oracle, 0x5EED, n, 3.0, nthreads=threads) == (out[2].sum, out[2].xor_)
    assert oracle_lib.triad_checksum(oracle, 0x5EED + 1, n, 3.0) != (out[2].sum, out[2].xor_)
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "triad_known_answers.json")))
    k = gold["answers"][str(64 << 20)]
    assert oracle_lib.triad_checksum(oracle, 0x5EED, (64 << 20) // 4, 3.0, nthreads=4) == (int(k["sum"], 16), int(k["xor"], 16))
