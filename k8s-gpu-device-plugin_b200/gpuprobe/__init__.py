"""ctypes binding of libgpuprobe.so (include/gpuprobe.h).

This is the harness-side twin of the cgo stub in INTEGRATION.md: tests,
``__graft_entry__.smoke()`` and ``bench.py`` drive the probe through exactly the
C ABI the Go plugin would bind.  There is no fallback of any kind: if the CUDA
library is missing, importing :func:`load` raises, and every call raises
:class:`GpuProbeError` on a non-zero return code.

Reference call sites this ABI serves (uppercaveman/k8s-gpu-device-plugin):
device/devices.go:74 (Health verdict), device/device_map.go:48-111 (dispatch),
plugin/plugin.go:173-189 (ListAndWatch health feed), router/api.go:27-37.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libgpuprobe.so")

GP_ABI_VERSION = 2

GP_PAT_WALKING_ONES, GP_PAT_WALKING_ZEROS, GP_PAT_ADDRESS = 0, 1, 2
GP_PROBE_COPY, GP_PROBE_TRIAD, GP_PROBE_PATTERN, GP_PROBE_GEMM = 1, 2, 4, 8
GP_PROBE_ALL_SINGLE = 0xF
GP_K_COPY, GP_K_TRIAD, GP_K_FILL, GP_K_VERIFY, GP_K_GEMM, GP_K_P2P, GP_K_INIT, GP_K_STEP = range(8)
GP_K_COUNT = 8
GP_K_NAMES = ["copy", "triad", "fill", "verify", "gemm", "p2p", "init", "step"]
GP_STREAM_AUTO, GP_STREAM_VEC, GP_STREAM_TMA = 0, 1, 2

GP_OK = 0
GP_ERR_INVALID, GP_ERR_NO_DEVICE, GP_ERR_CUDA, GP_ERR_NCCL = -1, -2, -3, -4
GP_ERR_NOMEM, GP_ERR_UNSUPPORTED, GP_ERR_STRUCT_SIZE, GP_ERR_TOO_LARGE = -5, -6, -7, -8
GP_ERR_DEVICE_LOST = -9
GP_ARENA_AUTO, GP_ARENA_VMM, GP_ARENA_MALLOC = 0, 1, 2
GP_GEMM_AUTO, GP_GEMM_1CTA_128x256, GP_GEMM_2CTA_256x256, GP_GEMM_2CTA_256x512 = 0, 1, 2, 3
GP_TEST_FAULT_GEMM_SKIP_TILE, GP_TEST_FAULT_GEMM_FLIP_BIT, GP_TEST_FAULT_P2P_CORRUPT = 1, 2, 4
GP_P2P_READ, GP_P2P_WRITE, GP_P2P_BIDIR_WRITE, GP_P2P_ALL_AT_ONCE = 0, 1, 2, 3
GP_P2P_BIDIR_READ, GP_P2P_COPY_ENGINE, GP_P2P_COPY_ENGINE_BIDIR = 4, 5, 6
GP_FABRIC_MAX_GPUS = 16


class GpuProbeError(RuntimeError):
    def __init__(self, code: int, what: str, detail: str):
        super().__init__(f"{what}: gp_error {code} ({detail})")
        self.code = code
        self.detail = detail


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("abi_version", C.c_uint32),
        ("hbm_bytes", C.c_uint64), ("pattern_bytes", C.c_uint64), ("arena_cap_bytes", C.c_uint64),
        ("gemm_mnk", C.c_uint32), ("warmup", C.c_uint32), ("iters", C.c_uint32),
        ("wo_passes", C.c_uint32), ("gemm_samples", C.c_uint32), ("stream_variant", C.c_uint32),
        ("keep_arena", C.c_uint32), ("gemm_output", C.c_uint32),
        ("seed", C.c_uint64),
        ("min_copy_gbps", C.c_double), ("min_triad_gbps", C.c_double),
        ("min_gemm_tflops", C.c_double), ("gemm_tol", C.c_double),
        ("min_p2p_gbps", C.c_double), ("max_p2p_spread", C.c_double),
        ("min_allreduce_busbw_gbps", C.c_double), ("fabric_bytes", C.c_uint64),
        ("arena_backend", C.c_uint32), ("sm_limit", C.c_uint32),
        ("gemm_variant", C.c_uint32), ("test_fault", C.c_uint32), ("reserved2", C.c_uint32 * 2),
        ("triad_expect_sum", C.c_uint64), ("triad_expect_xor", C.c_uint64),
    ]


class PatternResult(C.Structure):
    _fields_ = [("sum", C.c_uint64), ("xor_", C.c_uint64), ("wsum", C.c_uint64),
                ("mismatches", C.c_uint64), ("first_bad_index", C.c_uint64),
                ("first_bad_xor", C.c_uint64)]

    def astuple(self) -> Tuple[int, ...]:
        return (self.sum, self.xor_, self.wsum, self.mismatches, self.first_bad_index,
                self.first_bad_xor)


class Timing(C.Structure):
    _fields_ = [("best_ms", C.c_double), ("median_ms", C.c_double), ("mean_ms", C.c_double),
                ("launches", C.c_uint32), ("reserved", C.c_uint32)]


class GemmResult(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("m", C.c_uint32), ("n", C.c_uint32), ("k", C.c_uint32),
                ("t", Timing), ("tflops_best", C.c_double), ("tflops_median", C.c_double),
                ("max_scaled_err", C.c_double), ("freivalds_err", C.c_double),
                ("samples", C.c_uint32), ("samples_bad", C.c_uint32), ("c_checksum", C.c_uint64),
                ("output_bf16", C.c_uint32), ("reserved", C.c_uint32), ("bf16_mismatches", C.c_uint64)]


class DeviceResult(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("abi_version", C.c_uint32), ("uuid", C.c_char * 48),
        ("ordinal", C.c_int32), ("sm_count", C.c_int32), ("total_mem", C.c_uint64),
        ("cc_major", C.c_int32), ("cc_minor", C.c_int32),
        ("probed_mask", C.c_uint32), ("failed_mask", C.c_uint32),
        ("healthy", C.c_int32), ("error", C.c_int32), ("detail", C.c_int64),
        ("duration_s", C.c_double), ("hbm_bytes", C.c_uint64), ("pattern_bytes", C.c_uint64),
        ("copy_t", Timing), ("copy_gbps_best", C.c_double), ("copy_gbps_median", C.c_double),
        ("triad_t", Timing), ("triad_gbps_best", C.c_double), ("triad_gbps_median", C.c_double),
        ("fill_t", Timing), ("fill_gbps_best", C.c_double), ("fill_gbps_median", C.c_double),
        ("verify_t", Timing), ("verify_gbps_best", C.c_double), ("verify_gbps_median", C.c_double),
        ("copy_checksum", C.c_uint64), ("copy_ok", C.c_uint32), ("triad_ok", C.c_uint32),
        ("pattern_passes", C.c_uint32), ("pattern_bad_passes", C.c_uint32),
        ("pattern_mismatches", C.c_uint64),
        ("first_bad_kind", C.c_int32), ("first_bad_pass", C.c_int32),
        ("first_bad_index", C.c_uint64), ("first_bad_xor", C.c_uint64),
        ("pattern_digest", C.c_uint64), ("gemm", GemmResult),
        ("alloc_s", C.c_double), ("free_s", C.c_double), ("kernels_s", C.c_double),
        ("arena_bytes", C.c_uint64), ("arena_reused", C.c_uint32), ("arena_backend", C.c_uint32),
        ("triad_sum", C.c_uint64), ("triad_xor", C.c_uint64), ("triad_known_answer", C.c_uint32),
        ("free_deferred", C.c_uint32),
    ]


class FabricResult(C.Structure):
    _N2 = GP_FABRIC_MAX_GPUS * GP_FABRIC_MAX_GPUS
    _fields_ = [
        ("struct_size", C.c_uint32), ("ngpus", C.c_uint32), ("bytes", C.c_uint64),
        ("read_gbps", C.c_double * _N2), ("write_gbps", C.c_double * _N2),
        ("read_min", C.c_double), ("read_max", C.c_double), ("read_mean", C.c_double),
        ("write_min", C.c_double), ("write_max", C.c_double), ("write_mean", C.c_double),
        ("spread", C.c_double), ("worst_src", C.c_int32), ("worst_dst", C.c_int32),
        ("worst_gbps", C.c_double),
        ("allreduce_algbw_gbps", C.c_double), ("allreduce_busbw_gbps", C.c_double),
        ("allreduce_error", C.c_int32), ("data_ok", C.c_uint32), ("failed", C.c_uint32),
        ("healthy", C.c_int32), ("duration_s", C.c_double),
    ]

    def matrix(self, which: str) -> List[List[float]]:
        n = self.ngpus
        m = self.read_gbps if which == "read" else self.write_gbps
        return [[m[r * n + c] for c in range(n)] for r in range(n)]


class NvmlInfo(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("available", C.c_uint32), ("sm_clock_mhz", C.c_uint32),
                ("sm_clock_max_mhz", C.c_uint32), ("mem_clock_mhz", C.c_uint32), ("power_mw", C.c_uint32),
                ("temperature_c", C.c_uint32), ("other_compute_procs", C.c_uint32),
                ("throttle_reasons", C.c_uint64), ("mem_used_bytes", C.c_uint64), ("mem_total_bytes", C.c_uint64)]


class KernelStat(C.Structure):
    _fields_ = [("launches", C.c_uint64), ("total_ms", C.c_double), ("algo_bytes", C.c_double)]


# every symbol include/gpuprobe.h declares: (name, restype, argtypes)
_P = C.POINTER
ABI = [
    ("gp_init", C.c_int, [_P(Config), _P(C.c_void_p)]),
    ("gp_shutdown", None, [C.c_void_p]),
    ("gp_trim", C.c_int, [C.c_void_p]),
    ("gp_trim_device", C.c_int, [C.c_void_p, C.c_char_p]),
    ("gp_device_reset", C.c_int, [C.c_void_p, C.c_char_p]),
    ("gp_device_poisoned", C.c_int, [C.c_void_p, C.c_char_p]),
    ("gp_inject_cuda_error", C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
    ("gp_probe_fabric", C.c_int, [C.c_void_p, C.c_int, _P(FabricResult)]),
    ("gp_abi_version", C.c_uint32, []),
    ("gp_sizeof", C.c_size_t, [C.c_int]),
    ("gp_strerror", C.c_char_p, [C.c_int]),
    ("gp_last_error_detail", C.c_char_p, []),
    ("gp_device_count", C.c_int, [C.c_void_p]),
    ("gp_device_uuid", C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_size_t]),
    ("gp_nvml_snapshot", C.c_int, [C.c_void_p, C.c_char_p, _P(NvmlInfo)]),
    ("gp_probe_device", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint32, _P(DeviceResult)]),
    ("gp_probe_all", C.c_int, [C.c_void_p, C.c_int, C.c_uint32, _P(DeviceResult), _P(C.c_double)]),
    ("gp_pattern_fill", C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_uint64]),
    ("gp_pattern_verify", C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_uint64, _P(PatternResult)]),
    ("gp_pattern_host", C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, _P(PatternResult)]),
    ("gp_pattern_probe", C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_uint64, _P(PatternResult), _P(Timing), _P(Timing)]),
    ("gp_stream_probe", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint32, _P(Timing), _P(C.c_uint64)]),
    ("gp_stream_host", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_uint64]),
    ("gp_gemm_probe", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, _P(GemmResult)]),
    ("gp_gemm_host", C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]),
    ("gp_hbm_step", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint32, _P(PatternResult)]),
    ("gp_kernel_stats", C.c_int, [C.c_void_p, C.c_char_p, _P(KernelStat), C.c_int]),
    ("gp_inject_bitflip", C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.c_int]),
    ("gp_pattern_verify_noalloc", C.c_int, [C.c_int, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    ("gp_probe_p2p", C.c_int, [C.c_void_p, C.c_int, C.c_size_t, C.c_int, _P(C.c_double)]),
    ("gp_probe_allreduce", C.c_int, [C.c_void_p, C.c_int, C.c_size_t, C.c_uint32, C.c_uint32, _P(C.c_double), _P(C.c_double)]),
]

_lib = None


def load(path: Optional[str] = None) -> C.CDLL:
    """dlopen libgpuprobe.so and bind every declared symbol.  Raises if the
    library or any symbol is missing — there is no CPU stand-in."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(f"{p} is missing: build it with `make` (python -c 'import __graft_entry__ as g; g.build()')")
    lib = C.CDLL(p)
    for name, res, args in ABI:
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


GP_NOALLOC_MAX_WORDS = 1 << 35
GP_PATTERN_ALL_FREE = (1 << 64) - 1


def pattern_verify_noalloc(ordinal: int, d_words_ptr: int, nwords: int, kind: int, pas: int,
                           d_result_ptr: int, stream: int = 0) -> None:
    """gp_pattern_verify_noalloc: caller-owned device memory in, 48-byte device result out,
    enqueued on `stream` without synchronising and without allocating.  Needs no gp_ctx.
    Raises GpuProbeError on a non-zero return code."""
    lib = load()
    rc = lib.gp_pattern_verify_noalloc(ordinal, C.c_void_p(d_words_ptr), nwords, kind, pas,
                                       C.c_void_p(d_result_ptr), C.c_void_p(stream))
    if rc != GP_OK:
        raise GpuProbeError(rc, "gp_pattern_verify_noalloc",
                            (lib.gp_last_error_detail() or b"").decode(errors="replace"))


def _u(uuid) -> bytes:
    return uuid if isinstance(uuid, bytes) else str(uuid).encode()


class Probe:
    """Owns one gp_ctx.  Method names follow the C ABI without the gp_ prefix."""

    def __init__(self, **cfg):
        self.lib = load()
        c = Config()
        c.struct_size = C.sizeof(Config)
        c.abi_version = GP_ABI_VERSION
        for k, v in cfg.items():
            if not hasattr(c, k):
                raise TypeError(f"unknown gp_config field {k}")
            setattr(c, k, v)
        self.cfg = c
        self._ctx = C.c_void_p()
        self._check(self.lib.gp_init(C.byref(c), C.byref(self._ctx)), "gp_init")

    def _check(self, rc: int, what: str) -> None:
        if rc != GP_OK:
            raise GpuProbeError(rc, what, (self.lib.gp_last_error_detail() or b"").decode(errors="replace")
                                or (self.lib.gp_strerror(rc) or b"").decode())

    def close(self) -> None:
        if self._ctx:
            self.lib.gp_shutdown(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- enumeration
    def device_count(self) -> int:
        return self.lib.gp_device_count(self._ctx)

    def device_uuid(self, ordinal: int) -> str:
        buf = C.create_string_buffer(64)
        self._check(self.lib.gp_device_uuid(self._ctx, ordinal, buf, 64), "gp_device_uuid")
        return buf.value.decode()

    def trim(self) -> None:
        self._check(self.lib.gp_trim(self._ctx), "gp_trim")

    def trim_device(self, uuid) -> None:
        self._check(self.lib.gp_trim_device(self._ctx, _u(uuid)), "gp_trim_device")

    def device_reset(self, uuid) -> None:
        self._check(self.lib.gp_device_reset(self._ctx, _u(uuid)), "gp_device_reset")

    def device_poisoned(self, uuid) -> bool:
        rc = self.lib.gp_device_poisoned(self._ctx, _u(uuid))
        if rc < 0:
            self._check(rc, "gp_device_poisoned")
        return bool(rc)

    def inject_cuda_error(self, uuid, cuda_error: int) -> int:
        """Test-only: returns the gp_error the failing call would have returned (no exception)."""
        return self.lib.gp_inject_cuda_error(self._ctx, _u(uuid), cuda_error)

    def nvml_snapshot(self, uuid) -> NvmlInfo:
        out = NvmlInfo()
        self._check(self.lib.gp_nvml_snapshot(self._ctx, _u(uuid), C.byref(out)), "gp_nvml_snapshot")
        return out

    # -- verdicts
    def probe_device(self, uuid, mask: int = GP_PROBE_ALL_SINGLE) -> DeviceResult:
        out = DeviceResult()
        self._check(self.lib.gp_probe_device(self._ctx, _u(uuid), mask, C.byref(out)), "gp_probe_device")
        return out

    def probe_all(self, ngpus: int, mask: int = GP_PROBE_ALL_SINGLE) -> Tuple[List[DeviceResult], float]:
        arr = (DeviceResult * ngpus)()
        wall = C.c_double()
        self._check(self.lib.gp_probe_all(self._ctx, ngpus, mask, arr, C.byref(wall)), "gp_probe_all")
        return list(arr), wall.value

    # -- building blocks
    def pattern_fill(self, uuid, kind: int, pas: int, nwords: int) -> None:
        self._check(self.lib.gp_pattern_fill(self._ctx, _u(uuid), kind, pas, nwords), "gp_pattern_fill")

    def pattern_verify(self, uuid, kind: int, pas: int, nwords: int) -> PatternResult:
        out = PatternResult()
        self._check(self.lib.gp_pattern_verify(self._ctx, _u(uuid), kind, pas, nwords, C.byref(out)), "gp_pattern_verify")
        return out

    def pattern_host(self, uuid, words, kind: int, pas: int, fill: bool) -> PatternResult:
        """words: numpy uint64 array (C-contiguous) or None with fill=True."""
        out = PatternResult()
        ptr = C.c_void_p(words.ctypes.data) if words is not None and words.size else C.c_void_p()
        n = int(words.size) if words is not None else 0
        self._check(self.lib.gp_pattern_host(self._ctx, _u(uuid), ptr, n, kind, pas, int(fill), C.byref(out)), "gp_pattern_host")
        return out

    def pattern_probe(self, uuid, kind: int, pass_begin: int, pass_count: int, nwords: int):
        res = (PatternResult * pass_count)()
        ft, vt = Timing(), Timing()
        self._check(self.lib.gp_pattern_probe(self._ctx, _u(uuid), kind, pass_begin, pass_count, nwords, res, C.byref(ft), C.byref(vt)), "gp_pattern_probe")
        return list(res), ft, vt

    def stream_probe(self, uuid, which: int, nbytes: int, warmup: int = 2, iters: int = 10):
        t = Timing()
        chk = C.c_uint64()
        self._check(self.lib.gp_stream_probe(self._ctx, _u(uuid), which, nbytes, warmup, iters, C.byref(t), C.byref(chk)), "gp_stream_probe")
        return t, chk.value

    def stream_host(self, uuid, which: int, b, c, scalar: float, dst) -> None:
        nbytes = int(b.nbytes)
        cp = C.c_void_p(c.ctypes.data) if c is not None else C.c_void_p()
        self._check(self.lib.gp_stream_host(self._ctx, _u(uuid), which, C.c_void_p(b.ctypes.data), cp, scalar, C.c_void_p(dst.ctypes.data), nbytes), "gp_stream_host")

    def gemm_probe(self, uuid, mnk: int = 8192, seed: int = 0x5EED, warmup: int = 2, iters: int = 10, samples: int = 4096) -> GemmResult:
        out = GemmResult()
        self._check(self.lib.gp_gemm_probe(self._ctx, _u(uuid), mnk, seed, warmup, iters, samples, C.byref(out)), "gp_gemm_probe")
        return out

    def gemm_host(self, uuid, a, b, c, m: int, n: int, k: int) -> None:
        self._check(self.lib.gp_gemm_host(self._ctx, _u(uuid), C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_void_p(c.ctypes.data), m, n, k), "gp_gemm_host")

    def hbm_step(self, uuid, nbytes: int, step: int) -> List[PatternResult]:
        out = (PatternResult * 3)()
        self._check(self.lib.gp_hbm_step(self._ctx, _u(uuid), nbytes, step, out), "gp_hbm_step")
        return list(out)

    def kernel_stats(self, uuid, reset: bool = False) -> List[KernelStat]:
        out = (KernelStat * GP_K_COUNT)()
        self._check(self.lib.gp_kernel_stats(self._ctx, _u(uuid), out, int(reset)), "gp_kernel_stats")
        return list(out)

    def inject_bitflip(self, uuid, word_index: int, bit: int) -> None:
        self._check(self.lib.gp_inject_bitflip(self._ctx, _u(uuid), word_index, bit), "gp_inject_bitflip")

    # -- multi-GPU
    def probe_p2p(self, ngpus: int, nbytes: int, direction: int) -> List[List[float]]:
        arr = (C.c_double * (ngpus * ngpus))()
        self._check(self.lib.gp_probe_p2p(self._ctx, ngpus, nbytes, direction, arr), "gp_probe_p2p")
        return [[arr[r * ngpus + c] for c in range(ngpus)] for r in range(ngpus)]

    def probe_fabric(self, ngpus: int) -> FabricResult:
        out = FabricResult()
        self._check(self.lib.gp_probe_fabric(self._ctx, ngpus, C.byref(out)), "gp_probe_fabric")
        return out

    def probe_allreduce(self, nranks: int, nbytes: int, warmup: int = 2, iters: int = 10) -> Tuple[float, float]:
        alg, bus = C.c_double(), C.c_double()
        self._check(self.lib.gp_probe_allreduce(self._ctx, nranks, nbytes, warmup, iters, C.byref(alg), C.byref(bus)), "gp_probe_allreduce")
        return alg.value, bus.value
