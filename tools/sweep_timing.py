#!/usr/bin/env python3
"""Full-box health sweep timing: gp_probe_all on 1..N GPUs in one process (the plugin's shape),
cold arena (allocated inside; released on background threads as the call returns, and given time to arrive
before the next repetition: gp_trim) and warm arena.  Writes gpurun_out/sweep_<n>.json."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-device-plugin_b200"))
import gpuprobe
out = {}
t0 = time.time()
p = gpuprobe.Probe(keep_arena=0)
nmax = int(sys.argv[1]) if len(sys.argv) > 1 else p.device_count()
out["init_s"] = time.time() - t0
for n in [k for k in (1, 2, 4, 8) if k <= nmax]:
    rows = []
    for rep in range(3):
        res, wall = p.probe_all(n, gpuprobe.GP_PROBE_ALL_SINGLE)
        t1 = time.time()
        p.trim()
        rows.append({"wall_s": wall, "release_after_return_s": round(time.time() - t1, 4),
                     "alloc_wait_s": [round(r.alloc_s, 4) for r in res], "per_device_s": [round(r.duration_s, 3) for r in res], "healthy": [r.healthy for r in res],
                     "copy": [round(r.copy_gbps_median) for r in res], "triad": [round(r.triad_gbps_median) for r in res],
                     "fill": [round(r.fill_gbps_median) for r in res], "verify": [round(r.verify_gbps_median) for r in res],
                     "gemm": [round(r.gemm.tflops_median) for r in res]})
        print(n, rep, rows[-1], flush=True)
    out[str(n)] = rows
p.close()
p = gpuprobe.Probe(keep_arena=1)
for n in [k for k in (1, 2, 4, 8) if k <= nmax]:
    p.probe_all(n, gpuprobe.GP_PROBE_ALL_SINGLE)
    res, wall = p.probe_all(n, gpuprobe.GP_PROBE_ALL_SINGLE)
    out[f"{n}_warm_arena"] = {"wall_s": wall, "per_device_s": [round(r.duration_s, 3) for r in res]}
    print(n, "warm", out[f"{n}_warm_arena"], flush=True)
p.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"sweep_{nmax}.json"), "w"), indent=1)
