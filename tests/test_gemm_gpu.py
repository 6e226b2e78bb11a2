"""GPU parity: bf16 GEMM probe on tcgen05/TMEM through the C ABI vs the CPU oracle's fp64
reference of the same bf16-rounded inputs.

Stated tolerance: |c - ref| <= 1e-3 * SUM_k |a_ik| |b_jk| for the
fp32 output.  Products of two bf16 values are exact in fp32, so the only error is the fp32
accumulation over K; in practice it is ~1e-6 of SUM|a||b| and the tests also print it."""
import os

import numpy as np
import pytest

import gpuprobe
import oracle_lib

pytestmark = pytest.mark.gpu
DEV = "0"
TOL = 1e-3


VARIANTS = {"2cta_256x512": gpuprobe.GP_GEMM_AUTO, "2cta_256x256": gpuprobe.GP_GEMM_2CTA_256x256,
            "1cta_128x256": gpuprobe.GP_GEMM_1CTA_128x256}


@pytest.fixture(scope="module", params=list(VARIANTS), ids=list(VARIANTS))
def cta_group(request):
    """Every tcgen05 kernel the library has, selected through gp_config.gemm_variant."""
    return request.param


@pytest.fixture(scope="module")
def vprobe(cta_group):
    p = gpuprobe.Probe(keep_arena=1, hbm_bytes=1 << 30, gemm_variant=VARIANTS[cta_group])
    yield p
    p.close()


def ref_gemm(oracle, a, b, m, n, k):
    c = np.empty((m, n), dtype=np.float64)
    ad = np.empty((m, n), dtype=np.float64)
    oracle.gp_cpu_gemm_ref(a.ctypes.data, b.ctypes.data, c.ctypes.data, ad.ctypes.data, m, n, k)
    return c, ad


@pytest.mark.parametrize("m,n,k", [(256, 256, 64), (256, 512, 128), (512, 256, 1024),
                                   (768, 1024, 320), (256, 256, 8192), (2560, 2304, 192)])
def test_gemm_matches_oracle(vprobe, oracle, cta_group, m, n, k):
    a = oracle_lib.gen_bf16(oracle, 0x5EED + m, 0, m * k)
    b = oracle_lib.gen_bf16(oracle, (0x5EED + n) ^ oracle_lib.SEED_B_XOR, 0, n * k)
    c = np.zeros((m, n), dtype=np.float32)
    vprobe.gemm_host(DEV, a, b, c, m, n, k)
    ref, absdot = ref_gemm(oracle, a, b, m, n, k)
    assert np.isfinite(c).all()
    scaled = np.abs(c.astype(np.float64) - ref) / np.maximum(absdot, 1e-30)
    print(f"gemm {m}x{n}x{k} cta_group={cta_group}: max scaled err {scaled.max():.3e}")
    assert scaled.max() <= TOL


def test_gemm_layout_exact(vprobe, cta_group):
    """A selects one column of B per row: C[i][j] = B[j][i % K] exactly — any tile, swizzle or
    descriptor mistake shows up as a wrong element, not as rounding."""
    m, n, k = 512, 768, 256
    one = np.uint16(0x3F80)  # bf16 1.0
    a = np.zeros((m, k), dtype=np.uint16)
    a[np.arange(m), np.arange(m) % k] = one
    rng = np.random.default_rng(7)
    bf = rng.integers(-64, 64, size=(n, k)).astype(np.float32)  # small integers: exact in bf16
    b = (bf.view(np.uint32) >> 16).astype(np.uint16)
    c = np.zeros((m, n), dtype=np.float32)
    vprobe.gemm_host(DEV, a, b, c, m, n, k)
    expect = bf[:, np.arange(m) % k].T
    assert np.array_equal(c, expect)


def test_gemm_probe_small_and_full(cta_group):
    with gpuprobe.Probe(keep_arena=1, gemm_variant=VARIANTS[cta_group]) as p:
        for mnk, it in [(1024, 3), (2048, 3), (8192, 5)]:
            g = p.gemm_probe(DEV, mnk=mnk, iters=it, warmup=1, samples=4096)
            print(f"gemm probe {mnk}^3 cta_group={cta_group}: {g.tflops_best:.0f} TFLOP/s best, "
                  f"max scaled err {g.max_scaled_err:.2e}, freivalds {g.freivalds_err:.2e}")
            assert g.samples == 4096 and g.samples_bad == 0
            assert g.max_scaled_err <= TOL and g.freivalds_err <= TOL
            assert g.tflops_best > 1.0
        # same seed -> same checksum (deterministic accumulation order)
        g2 = p.gemm_probe(DEV, mnk=2048, iters=1, warmup=0, samples=16)
        g3 = p.gemm_probe(DEV, mnk=2048, iters=1, warmup=0, samples=16)
        assert g2.c_checksum == g3.c_checksum


def test_gemm_detects_corruption(probe, oracle):
    """The host-side sampled check must actually be able to fail: a different seed for the
    reference than for the operands is a stand-in for a corrupted tensor pipe."""
    m = n = k = 256
    a = oracle_lib.gen_bf16(oracle, 1, 0, m * k)
    b = oracle_lib.gen_bf16(oracle, 2, 0, n * k)
    c = np.zeros((m, n), dtype=np.float32)
    probe.gemm_host(DEV, a, b, c, m, n, k)
    a2 = a.copy()
    a2[5 * k + 3] ^= np.uint16(0x4000)  # flip one exponent bit of A[5][3]
    ref, absdot = ref_gemm(oracle, a2, b, m, n, k)
    scaled = np.abs(c.astype(np.float64) - ref) / np.maximum(absdot, 1e-30)
    assert scaled[5].max() > TOL and np.delete(scaled, 5, axis=0).max() <= TOL


def test_gemm_invalid_shapes(probe):
    z = np.zeros(256 * 256, dtype=np.uint16)
    c = np.zeros((256, 256), dtype=np.float32)
    for m, n, k in [(128, 256, 64), (256, 255, 64), (256, 256, 32), (0, 256, 64)]:
        with pytest.raises(gpuprobe.GpuProbeError) as e:
            probe.gemm_host(DEV, z, z, c, m, n, k)
        assert e.value.code == gpuprobe.GP_ERR_INVALID


def test_gemm_golden_points(vprobe, oracle, cta_group):
    """The generated problem at the committed fixture positions (tests/golden/stream_gemm_golden.json)."""
    import json
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "stream_gemm_golden.json")))
    for g in gold["gemm"]:
        n = g["mnk"]
        a = oracle_lib.gen_bf16(oracle, g["seed"], 0, n * n)
        b = oracle_lib.gen_bf16(oracle, g["seed"] ^ oracle_lib.SEED_B_XOR, 0, n * n)
        c = np.zeros((n, n), dtype=np.float32)
        vprobe.gemm_host(DEV, a, b, c, n, n, n)
        for pt in g["points"]:
            assert abs(float(c[pt["i"], pt["j"]]) - pt["ref"]) <= TOL * pt["absdot"], pt


def test_gemm_probe_bf16_output(cta_group):
    """C in bf16 (gp_config.gemm_output = 1): the timed launches write bf16; the strict fp64 checks run on
    the fp32 C of the same launch plan and the bf16 C must equal RNE(bf16) of it BIT FOR BIT (the
    accumulators are the same bits), so there is no rounding allowance anywhere."""
    with gpuprobe.Probe(keep_arena=1, gemm_output=1, gemm_variant=VARIANTS[cta_group]) as p:
        for mnk in (1024, 2048, 8192):
            g = p.gemm_probe(DEV, mnk=mnk, iters=3, warmup=1, samples=4096)
            print(f"gemm probe bf16-out {mnk}^3 {cta_group}: {g.tflops_best:.0f} TFLOP/s best, "
                  f"max scaled err {g.max_scaled_err:.2e}, freivalds {g.freivalds_err:.2e}")
            assert g.output_bf16 == 1 and g.samples_bad == 0 and g.bf16_mismatches == 0
            assert g.max_scaled_err <= TOL and g.freivalds_err <= TOL
    with gpuprobe.Probe(keep_arena=1) as p:
        assert p.gemm_probe(DEV, mnk=1024, iters=1, warmup=0, samples=64).output_bf16 == 0


@pytest.mark.parametrize("out", [0, 1], ids=["fp32_out", "bf16_out"])
def test_gemm_check_sees_an_unwritten_tile(out):
    """C is pre-filled with NaN bit patterns so that a tile the kernel never stored shows.  The checks
    must then FAIL: a NaN sample is an error outright (std::max and <= silently drop NaN — round 1's
    check graded such a result healthy), and Freivalds sees the NaN row."""
    with gpuprobe.Probe(keep_arena=1, gemm_output=out, test_fault=gpuprobe.GP_TEST_FAULT_GEMM_SKIP_TILE) as p:
        g = p.gemm_probe(DEV, mnk=2048, iters=1, warmup=0, samples=4096)
        assert g.samples_bad > 0
        if out == 0:
            assert g.max_scaled_err >= 1e29 and g.freivalds_err >= 1e29   # both checks flagged non-finite data
        else:
            assert g.bf16_mismatches == 256 * 256                        # the bf16 tile no longer equals RNE(fp32)
        r = p.probe_device(DEV, gpuprobe.GP_PROBE_GEMM)
        assert r.healthy == 0 and r.failed_mask == gpuprobe.GP_PROBE_GEMM
    with gpuprobe.Probe(keep_arena=1, gemm_output=out) as p:  # and the same problem without the fault is clean
        assert p.gemm_probe(DEV, mnk=2048, iters=1, warmup=0, samples=4096).samples_bad == 0


def test_gemm_bf16_output_is_checked_bit_exact():
    """One flipped low mantissa bit in ONE element of the bf16 C: far inside any tolerance, still caught."""
    with gpuprobe.Probe(keep_arena=1, gemm_output=1, test_fault=gpuprobe.GP_TEST_FAULT_GEMM_FLIP_BIT) as p:
        g = p.gemm_probe(DEV, mnk=1024, iters=1, warmup=0, samples=1024)
        assert g.bf16_mismatches == 1 and g.samples_bad == 1
        assert g.max_scaled_err <= TOL and g.freivalds_err <= TOL   # the fp32 run itself is clean


def test_gemm_reference_cache_is_keyed_by_seed_shape_and_samples():
    """The host fp64 reference of the sampled entries is computed once per context and reused; a different seed,
    shape or sample count must get its own (a stale reference would show as thousands of bad samples)."""
    with gpuprobe.Probe(keep_arena=1) as p:
        for seed, mnk, samples in [(1, 1024, 512), (2, 1024, 512), (1, 1024, 512), (1, 2048, 512), (1, 1024, 1024),
                                   (2, 1024, 512), (3, 512, 64), (4, 512, 64), (5, 512, 64), (6, 512, 64),
                                   (7, 512, 64), (8, 512, 64), (9, 512, 64), (1, 1024, 512)]:  # > 8 keys: eviction too
            g = p.gemm_probe(DEV, mnk=mnk, seed=seed, iters=1, warmup=0, samples=samples)
            assert g.samples == samples and g.samples_bad == 0 and g.max_scaled_err <= TOL, (seed, mnk, samples)
