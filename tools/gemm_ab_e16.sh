cd /root/repo
E16=$PWD/k8s-gpu-device-plugin_b200/libgpuprobe_e16.so
for i in 1 2 3; do
  python tools/gemm_only.py 8192 20
  GPUPROBE_LIB=$E16 python tools/gemm_only.py 8192 20
done
GEMM_OUT_BF16=1 python tools/gemm_only.py 8192 20
GEMM_OUT_BF16=1 GPUPROBE_LIB=$E16 python tools/gemm_only.py 8192 20
GEMM_VARIANT=2 python tools/gemm_only.py 8192 20
GEMM_VARIANT=2 GPUPROBE_LIB=$E16 python tools/gemm_only.py 8192 20
