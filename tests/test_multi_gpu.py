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
    for direction in (0, 1, 2, 4):
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
