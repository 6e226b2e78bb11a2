#!/usr/bin/env python3
"""Runs only the GEMM probe (for ncu captures): python tools/gemm_only.py [mnk] [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-device-plugin_b200"))
import gpuprobe
if os.environ.get("GPUPROBE_LIB"):  # e.g. the development build with the environment switches (make dev)
    gpuprobe._lib = gpuprobe.load(os.environ["GPUPROBE_LIB"])
mnk = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
out = int(os.environ.get("GEMM_OUT_BF16", "0"))
variant = int(os.environ.get("GEMM_VARIANT", "0"))
with gpuprobe.Probe(keep_arena=1, gemm_output=out, gemm_variant=variant) as p:
    g = p.gemm_probe("0", mnk=mnk, iters=iters, warmup=2, samples=1024)
    print(f"gemm {mnk}^3 variant={variant} group_m={os.environ.get('GPUPROBE_GEMM_GROUP_M', '-')} hints={os.environ.get('GPUPROBE_GEMM_HINTS', '-')} out={'bf16' if out else 'fp32'}: best {g.tflops_best:.1f} median {g.tflops_median:.1f} TFLOP/s, best {g.t.best_ms:.4f} ms, "
          f"err {g.max_scaled_err:.2e}, freivalds {g.freivalds_err:.2e}, bad {g.samples_bad}")
