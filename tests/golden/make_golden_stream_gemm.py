"""Generates tests/golden/stream_gemm_golden.json from the CPU oracle (the reference has no
vectors for this path — see make_golden.py).  Triad: checksum of a = b + 3c over the generator's
operands.  GEMM: fp64 reference values and SUM|a||b| at fixed positions of the generated problem.

    python tests/golden/make_golden_stream_gemm.py
"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib  # noqa: E402

lib = oracle_lib.load()
out = {"source": "oracle/gpuprobe_oracle.c (parity unpinned: the reference holds no vectors)", "triad": [], "gemm": []}
for nbytes in (1 << 20, 8 << 20):
    n = nbytes // 4
    a = np.empty(n, dtype=np.float32)
    lib.gp_cpu_triad_gen(a.ctypes.data, 0x5EED, 3.0, 0, n)
    r = oracle_lib.verify(lib, a.view(np.uint64), 0, 0)
    out["triad"].append({"nbytes": nbytes, "seed": 0x5EED, "scalar": 3.0, "sum": hex(r.sum), "xor": hex(r.xor_), "wsum": hex(r.wsum)})
for mnk in (256, 1024):
    pts = []
    for (i, j) in [(0, 0), (1, 2), (mnk - 1, mnk - 1), (127, 128), (128, 127), (mnk // 2, 5), (37, mnk - 3), (255, 0)]:
        ad = C.c_double()
        v = lib.gp_cpu_gemm_sample(0x5EED, mnk, i, j, C.byref(ad))
        pts.append({"i": i, "j": j, "ref": v, "absdot": ad.value})
    out["gemm"].append({"mnk": mnk, "seed": 0x5EED, "points": pts})
with open(os.path.join(HERE, "stream_gemm_golden.json"), "w") as f:
    json.dump(out, f, indent=1)
print("ok")
