"""C-ABI conformance without a GPU: the library loads, exports every symbol that
include/gpuprobe.h declares, agrees with the binding on struct layout, and fails
LOUDLY (no CPU stand-in) when no CUDA device is present."""
import ctypes as C
import os
import re
import subprocess

import pytest

import gpuprobe

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gpuprobe.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = gpuprobe.load()
    names = declared_symbols()
    assert len(names) >= 24
    for n in names:
        assert hasattr(lib, n), f"libgpuprobe.so does not export {n}"
    bound = {n for n, _, _ in gpuprobe.ABI}
    assert bound == set(names), (bound ^ set(names))


def test_struct_layout_matches_library():
    lib = gpuprobe.load()
    assert lib.gp_abi_version() == gpuprobe.GP_ABI_VERSION
    for which, cls in enumerate([gpuprobe.Config, gpuprobe.PatternResult, gpuprobe.Timing,
                                 gpuprobe.GemmResult, gpuprobe.DeviceResult, gpuprobe.KernelStat,
                                 gpuprobe.NvmlInfo, gpuprobe.FabricResult]):
        assert lib.gp_sizeof(which) == C.sizeof(cls), cls.__name__
    assert lib.gp_sizeof(99) == 0
    assert C.sizeof(gpuprobe.PatternResult) == 48  # "result struct 48 B per pass"


def test_strerror_and_bad_config():
    lib = gpuprobe.load()
    assert lib.gp_strerror(0) == b"ok"
    assert b"CUDA" in lib.gp_strerror(gpuprobe.GP_ERR_CUDA)
    cfg = gpuprobe.Config()
    cfg.struct_size = 8  # skewed caller
    cfg.abi_version = gpuprobe.GP_ABI_VERSION
    ctx = C.c_void_p()
    assert lib.gp_init(C.byref(cfg), C.byref(ctx)) == gpuprobe.GP_ERR_STRUCT_SIZE
    assert not ctx.value
    assert lib.gp_init(None, None) == gpuprobe.GP_ERR_INVALID


def _has_gpu():
    try:
        return subprocess.run(["nvidia-smi", "-L"], capture_output=True, timeout=20).returncode == 0
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly_not_silently():
    with pytest.raises(gpuprobe.GpuProbeError) as e:
        gpuprobe.Probe()
    assert e.value.code in (gpuprobe.GP_ERR_CUDA, gpuprobe.GP_ERR_NO_DEVICE)


def test_product_does_not_reference_oracle():
    """The product tree must never import, link or name the oracle."""
    pkg = os.path.join(ROOT, "k8s-gpu-device-plugin_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".hpp", ".cpp", ".py", ".go")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "gp_cpu_" not in txt, f
                assert "libgpuprobe_oracle" not in txt and "oracle_lib" not in txt, f
    out = subprocess.run(["ldd", gpuprobe.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out


def _build_c_driver(tmp_path):
    exe = str(tmp_path / "c_abi_driver")
    pkg = os.path.join(ROOT, "k8s-gpu-device-plugin_b200")
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c_abi_driver.c"), "-o", exe, "-L", pkg, "-lgpuprobe",
                           "-Wl,-rpath," + pkg])
    return exe


@pytest.mark.skipif(_has_gpu(), reason="CPU variant; the GPU variant is test_c_driver_on_gpu")
def test_header_is_valid_c99_and_driver_fails_loudly_without_gpu(tmp_path):
    """cgo compiles the header as C: it must be valid C99, and a C caller must see the same
    struct sizes as the library."""
    r = subprocess.run([_build_c_driver(tmp_path)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 3, (r.returncode, r.stdout, r.stderr)
    assert "no GPU: gp_init" in r.stdout


@pytest.mark.gpu
def test_c_driver_on_gpu(tmp_path):
    r = subprocess.run([_build_c_driver(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "c_abi_driver ok" in r.stdout
