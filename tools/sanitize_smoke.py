#!/usr/bin/env python3
"""Small pass over every kernel for compute-sanitizer (memcheck): no oracle, tiny sizes."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-device-plugin_b200"))
import gpuprobe
for variant in (gpuprobe.GP_STREAM_TMA, gpuprobe.GP_STREAM_VEC):
    with gpuprobe.Probe(keep_arena=1, hbm_bytes=8 << 20, pattern_bytes=4 << 20, wo_passes=2, iters=1, warmup=1,
                        gemm_mnk=512, gemm_samples=64, stream_variant=variant) as p:
        r = p.probe_device("0", gpuprobe.GP_PROBE_ALL_SINGLE)
        assert r.healthy == 1, (r.failed_mask, r.error)
        out = p.hbm_step("0", 1 << 20, 3)
        assert out[0].mismatches == 0 and out[1].mismatches == 0
        n = 64 * 33
        w = np.arange(n, dtype=np.uint64)
        p.pattern_host("0", w, 2, 0, fill=False)
        b = np.ones(32 * 1027 // 4, dtype=np.float32); a = np.zeros_like(b)
        p.stream_host("0", gpuprobe.GP_PROBE_TRIAD, b, b, 3.0, a)
        assert np.all(a == 4.0)
        m = p.probe_p2p(1, 1 << 20, 1)
print("sanitize_smoke ok")
