"""The N>1 plumbing of bench.py on CPU (gloo, world size 2): the reference arm under torchrun
prints exactly one line from rank 0 and the other rank exits 0 without work; the
max-over-ranks helper returns the maximum on every rank."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_under_torchrun_world2():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "bench.py"),
                        "--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1",
                        "--cpu-bytes", str(8 << 20)], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1  # rank 0 alone prints
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["unit"] == "GB/s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import bench
    dist.init_process_group("gloo", rank=rank, world_size=world)
    got = bench.max_over_ranks(dist, [1.0 + rank, 10.0 - rank], "cpu")
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, got))


def test_max_over_ranks_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29534, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out[0] == out[1] == [2.0, 10.0]


def test_max_over_ranks_single_process_is_identity():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.max_over_ranks(None, [3.0, 4.0], "cpu") == [3.0, 4.0]
    wo, ad = bench.closed_forms(1 << 31, 0)
    assert wo == (0xFFFFFFFFFE000000, 0) and ad == (0xFFFFFFFE00000000, 0)
