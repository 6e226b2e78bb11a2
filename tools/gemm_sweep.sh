#!/bin/bash
# GEMM raster / variant sweep with the development build (make dev): one process per setting, because the
# switches are read once.  Usage: tools/gemm_sweep.sh > gpurun_out/gemm_sweep.txt
cd "$(dirname "$0")/.."
export GPUPROBE_LIB=$PWD/k8s-gpu-device-plugin_b200/libgpuprobe_dev.so
for v in 0 2 1; do GEMM_VARIANT=$v python tools/gemm_only.py 8192 20; done
for g in 4 8 16 32; do GPUPROBE_GEMM_GROUP_M=$g python tools/gemm_only.py 8192 20; done
for h in 0 5 37 21 38; do GPUPROBE_GEMM_HINTS=$h python tools/gemm_only.py 8192 20; done
GEMM_OUT_BF16=1 python tools/gemm_only.py 8192 20
for g in 4 8 16; do GEMM_VARIANT=2 GPUPROBE_GEMM_GROUP_M=$g python tools/gemm_only.py 8192 20; done
