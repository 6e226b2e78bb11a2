#!/usr/bin/env python3
"""Runs the cross-device probes over the size ladder of SURVEY.md §8 (64 KiB .. 1 GiB x4 for
P2P; the same plus 4 GiB for all-reduce) and writes gpurun_out/multi_gpu_<n>.json.
Usage: python tools/multi_gpu_report.py [ngpus]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-device-plugin_b200"))
import gpuprobe  # noqa: E402

p = gpuprobe.Probe(keep_arena=1, arena_cap_bytes=8 << 30)
n = int(sys.argv[1]) if len(sys.argv) > 1 else p.device_count()
out = {"n_gpus": n, "p2p": {}, "allreduce": {}}
sizes = [64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30]
for name, d in (("read", 0), ("write", 1), ("bidir_write", 2), ("bidir_read", 4), ("all_at_once_read", 3)):
    out["p2p"][name] = {}
    for s in sizes:
        m = p.probe_p2p(n, s, d)
        off = [m[i][j] for i in range(n) for j in range(n) if i != j]
        out["p2p"][name][str(s)] = {"min": min(off) if off else None, "max": max(off) if off else None,
                                    "mean": sum(off) / len(off) if off else None,
                                    "local": sum(m[i][i] for i in range(n)) / n,
                                    "matrix": m if s >= (256 << 20) else None}
        print(name, s, out["p2p"][name][str(s)]["min"], out["p2p"][name][str(s)]["max"], flush=True)
for k in sorted({2, 4, 8, n}):
    if k > n or k < 2:
        continue
    out["allreduce"][str(k)] = {}
    t0 = time.time()
    for s in sizes + [4 << 30]:
        alg, bus = p.probe_allreduce(k, s, warmup=2, iters=10 if s <= (64 << 20) else 5)
        out["allreduce"][str(k)][str(s)] = {"algbw_gbps": alg, "busbw_gbps": bus}
        print("allreduce", k, s, round(alg, 1), round(bus, 1), flush=True)
    out["allreduce"][str(k)]["seconds_including_comm_init"] = time.time() - t0
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", f"multi_gpu_{n}.json"), "w") as f:
    json.dump(out, f, indent=1)
p.close()
