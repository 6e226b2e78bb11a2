"""ctypes loader for oracle/libgpuprobe_oracle.so — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs import this module.  The oracle is "parity unpinned": the reference holds no
implementation or golden vector for this path (see oracle/gpuprobe_oracle.c).
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libgpuprobe_oracle.so")


class PatternResult(C.Structure):
    _fields_ = [("sum", C.c_uint64), ("xor_", C.c_uint64), ("wsum", C.c_uint64),
                ("mismatches", C.c_uint64), ("first_bad_index", C.c_uint64),
                ("first_bad_xor", C.c_uint64)]

    def astuple(self):
        return (self.sum, self.xor_, self.wsum, self.mismatches, self.first_bad_index,
                self.first_bad_xor)


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", ORACLE_DIR])
    lib = C.CDLL(ORACLE_SO)
    P = C.POINTER
    lib.gp_cpu_pattern.restype = C.c_int
    lib.gp_cpu_pattern.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, P(PatternResult)]
    lib.gp_cpu_pattern_fill.restype = C.c_int
    lib.gp_cpu_pattern_fill.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int]
    lib.gp_cpu_pattern_verify.restype = C.c_int
    lib.gp_cpu_pattern_verify.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, P(PatternResult)]
    lib.gp_cpu_pattern_word.restype = C.c_uint64
    lib.gp_cpu_pattern_word.argtypes = [C.c_int, C.c_int, C.c_uint64]
    lib.gp_cpu_digest_step.restype = C.c_uint64
    lib.gp_cpu_digest_step.argtypes = [C.c_uint64, P(PatternResult), C.c_uint32]
    lib.gp_cpu_mix64.restype = C.c_uint64
    lib.gp_cpu_mix64.argtypes = [C.c_uint64, C.c_uint64]
    lib.gp_cpu_gen_f32.restype = C.c_float
    lib.gp_cpu_gen_f32.argtypes = [C.c_uint64, C.c_uint64]
    lib.gp_cpu_gen_bf16.restype = C.c_uint16
    lib.gp_cpu_gen_bf16.argtypes = [C.c_uint64, C.c_uint64]
    lib.gp_cpu_gen_bf16_array.restype = None
    lib.gp_cpu_gen_bf16_array.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
    lib.gp_cpu_gen_f32_array.restype = None
    lib.gp_cpu_gen_f32_array.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
    lib.gp_cpu_triad.restype = None
    lib.gp_cpu_triad.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_uint64]
    lib.gp_cpu_triad_gen.restype = None
    lib.gp_cpu_triad_gen.argtypes = [C.c_void_p, C.c_uint64, C.c_float, C.c_uint64, C.c_uint64]
    lib.gp_cpu_sum_words.restype = C.c_uint64
    lib.gp_cpu_sum_words.argtypes = [C.c_void_p, C.c_uint64]
    lib.gp_cpu_gemm_ref.restype = None
    lib.gp_cpu_gemm_ref.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
    lib.gp_cpu_gemm_sample.restype = C.c_double
    lib.gp_cpu_gemm_sample.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, P(C.c_double)]
    lib.gp_cpu_triad_checksum.restype = None
    lib.gp_cpu_triad_checksum.argtypes = [C.c_uint64, C.c_uint64, C.c_float, C.c_int, P(C.c_uint64), P(C.c_uint64)]
    lib.gp_cpu_hbm_step.restype = C.c_int
    lib.gp_cpu_hbm_step.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint64, C.c_int, C.c_int, P(PatternResult)]
    _lib = lib
    return lib


SEED_B_XOR = 0xA5A5A5A5A5A5A5A5
SEED_TRIAD_B_XOR = 0x1111111111111111
SEED_TRIAD_C_XOR = 0x2222222222222222


def pattern(lib, nwords, kind, pas, nthreads=1, buf=None):
    """fill + verify; returns (PatternResult, buffer)."""
    if buf is None:
        buf = np.empty(nwords, dtype=np.uint64)
    r = PatternResult()
    rc = lib.gp_cpu_pattern(buf.ctypes.data, nwords, kind, pas, nthreads, C.byref(r))
    assert rc == 0, rc
    return r, buf


def verify(lib, buf, kind, pas, nthreads=1):
    r = PatternResult()
    rc = lib.gp_cpu_pattern_verify(buf.ctypes.data, buf.size, kind, pas, nthreads, C.byref(r))
    assert rc == 0, rc
    return r


def gen_f32(lib, seed, idx0, count):
    out = np.empty(count, dtype=np.float32)
    lib.gp_cpu_gen_f32_array(seed, idx0, count, out.ctypes.data)
    return out


def gen_bf16(lib, seed, idx0, count):
    out = np.empty(count, dtype=np.uint16)
    lib.gp_cpu_gen_bf16_array(seed, idx0, count, out.ctypes.data)
    return out


def hbm_step(lib, arena, nbytes, step, seed=0x5EED, gen=True, nthreads=1):
    """The CPU port of gp_hbm_step over a 5*nbytes numpy uint8 arena; returns 3 results."""
    out = (PatternResult * 3)()
    rc = lib.gp_cpu_hbm_step(arena.ctypes.data, nbytes, step, seed, int(gen), nthreads, out)
    assert rc == 0, rc
    return list(out)


def triad_checksum(lib, seed, nfloats, s=3.0, nthreads=1):
    """(sum, xor) of the u64 words of a = fmaf(s, c, b) over the generator's b and c — the triad probe's known answer."""
    a, b = C.c_uint64(), C.c_uint64()
    lib.gp_cpu_triad_checksum(seed, nfloats, s, nthreads, C.byref(a), C.byref(b))
    return a.value, b.value
