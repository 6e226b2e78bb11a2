"""Cross-device probes through the C ABI: in-kernel NVLink P2P matrix and NCCL all-reduce
bus bandwidth.  With one visible GPU only the n=1 degenerate cases run (local bandwidth on the
diagonal, a one-rank all-reduce); the pair tests need `gpurun --gpus 2` or more."""
import pytest

import gpuprobe

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mp():
    # libgpuprobe dlopens "libnccl.so.2" and reuses whichever copy the process already holds.
    # torch bundles a newer NCCL under the same soname; if the system copy were loaded first, a
    # later `import torch` in this pytest process would bind to it and fail on a missing symbol.
    # Importing torch first makes its copy the process-wide one (bench.py does the same).
    import torch  # noqa: F401
    p = gpuprobe.Probe(keep_arena=1)
    yield p
    p.close()


def test_rounds_cover_every_pair_once():
    # host logic of the schedule (mirrors rounds_for in csrc/p2p.cu): circle method
    def rounds_for(n):
        m = n + (n % 2)
        out = []
        for r in range(m - 1):
            rd = []
            for i in range(m // 2):
                a = (r + i) % (m - 1)
                b = m - 1 if i == 0 else (r + m - 1 - i) % (m - 1)
                if a < n and b < n:
                    rd.append((min(a, b), max(a, b)))
            out.append(rd)
        return out
    for n in (2, 3, 4, 5, 8):
        seen = set()
        for rd in rounds_for(n):
            used = [x for pr in rd for x in pr]
            assert len(used) == len(set(used))  # no GPU serves two peers in a round
            for pr in rd:
                assert pr not in seen
                seen.add(pr)
        assert len(seen) == n * (n - 1) // 2


def test_p2p_single_gpu_diagonal(mp):
    m = mp.probe_p2p(1, 256 << 20, 0)
    assert m[0][0] > 500  # local read bandwidth, GB/s
    m = mp.probe_p2p(1, 256 << 20, 1)
    assert m[0][0] > 500


def test_allreduce_single_rank(mp):
    alg, bus = mp.probe_allreduce(1, 64 << 20, warmup=1, iters=3)
    assert alg > 0 and bus == 0.0  # 2(n-1)/n = 0 for n = 1


def test_p2p_matrix_pairs(mp):
    n = mp.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    for direction in (0, 1, 2, 4, 5, 6):  # 5, 6: the cudaMemcpyPeerAsync yardstick, same schedule
        m = mp.probe_p2p(n, 256 << 20, direction)
        off = [m[i][j] for i in range(n) for j in range(n) if i != j]
        assert min(off) > 100, (direction, m)          # NVLink, not PCIe
        assert max(off) < 1000, (direction, m)         # not above the 900 GB/s/direction link
        assert min(off) > 0.6 * max(off), (direction, m)  # NVSwitch: flat off-diagonal
    m = mp.probe_p2p(n, 256 << 20, 3)  # all pairs at once: aggregate ingress per GPU on the diagonal
    for i in range(n):
        assert 100 < m[i][i] < 1000 and abs(m[i][(i + 1) % n] * (n - 1) - m[i][i]) < 1e-6 * m[i][i]
    with pytest.raises(gpuprobe.GpuProbeError):
        mp.probe_p2p(n + 1, 1 << 20, 0)


def test_allreduce_pairs(mp):
    n = mp.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    alg, bus = mp.probe_allreduce(n, 256 << 20, warmup=2, iters=5)
    assert abs(bus - alg * 2 * (n - 1) / n) < 1e-6 * max(bus, 1)
    assert bus > 50


def test_p2p_every_mode_checks_its_data_on_one_gpu(mp):
    """The diagonal runs the same kernels on local memory; every read folds what it loaded and the host
    compares it with the closed form of the address pattern (a read that returned garbage at full speed
    would otherwise grade healthy), every write and copy is verified where it landed."""
    for direction in (gpuprobe.GP_P2P_READ, gpuprobe.GP_P2P_WRITE, gpuprobe.GP_P2P_COPY_ENGINE):
        for nbytes in (64 << 10, 1 << 20, (16 << 20) + 32):   # the last one is ragged for every launch shape
            m = mp.probe_p2p(1, nbytes, direction)
            assert m[0][0] > 0


def test_fabric_verdict_single_gpu_and_floors(mp):
    f = mp.probe_fabric(1)
    assert f.healthy == 1 and f.data_ok == 1 and f.failed == 0 and f.ngpus == 1
    assert f.read_gbps[0] > 500 and f.write_gbps[0] > 500 and f.worst_src == -1
    with pytest.raises(gpuprobe.GpuProbeError):
        mp.probe_fabric(gpuprobe.GP_FABRIC_MAX_GPUS + 1)


def test_fabric_verdict_pairs():
    """On >= 2 GPUs: flat NVSwitch matrix, floors enforced in both directions of the comparison."""
    import torch  # noqa: F401  (NCCL soname, see the fixture above)
    with gpuprobe.Probe(keep_arena=1, fabric_bytes=256 << 20) as p:
        n = p.device_count()
        if n < 2:
            pytest.skip("needs >= 2 GPUs")
        f = p.probe_fabric(n)
        assert f.healthy == 1 and f.data_ok == 1
        assert 100 < f.read_min <= f.read_max < 1000 and 100 < f.write_min <= f.write_max < 1000
        assert 1.0 <= f.spread < 1.5 and f.worst_gbps == min(f.read_min, f.write_min)
        assert f.allreduce_error == 0 and f.allreduce_busbw_gbps > 50
    # a floor no fabric can meet: the same measurement must now grade unhealthy, with the reason bits set
    with gpuprobe.Probe(keep_arena=1, fabric_bytes=64 << 20, min_p2p_gbps=5000.0, max_p2p_spread=1.0000001,
                        min_allreduce_busbw_gbps=1e6) as p:
        f = p.probe_fabric(n)
        assert f.healthy == 0 and (f.failed & 1) and (f.failed & 4) and f.data_ok == 1


def test_a_read_that_returns_wrong_bytes_is_noticed():
    """One flipped bit in 64 MiB of a source buffer (test hook): the read that loads it folds a different sum and
    xor than the closed form of the pattern, so gp_probe_p2p fails instead of reporting a bandwidth, the copy-engine
    mode finds the word where it landed, and the fabric verdict says data_ok = 0 / unhealthy — on however many GPUs
    there are (with one, the reader is the GPU itself)."""
    import torch  # noqa: F401
    with gpuprobe.Probe(keep_arena=1, fabric_bytes=64 << 20, test_fault=gpuprobe.GP_TEST_FAULT_P2P_CORRUPT) as p:
        n = p.device_count()
        for mode in (gpuprobe.GP_P2P_READ, gpuprobe.GP_P2P_COPY_ENGINE):
            with pytest.raises(gpuprobe.GpuProbeError) as e:
                p.probe_p2p(n, 64 << 20, mode)
            assert e.value.code == gpuprobe.GP_ERR_CUDA and ("wrong bytes" in e.value.detail or "did not land intact" in e.value.detail)
        f = p.probe_fabric(n)
        assert f.data_ok == 0 and f.healthy == 0 and (f.failed & 8)
    with gpuprobe.Probe(keep_arena=1, fabric_bytes=64 << 20) as p:   # the same without the fault is clean
        f = p.probe_fabric(p.device_count())
        assert f.data_ok == 1 and f.healthy == 1
