"""Generates tests/golden/pattern_golden.json.

The reference (uppercaveman/k8s-gpu-device-plugin) has no implementation, test or
fixture for the probe path, so these vectors cannot come from it; they are produced
by this repo's oracle (oracle/gpuprobe_oracle.c) after test_oracle.py has pinned it
to the closed forms of SURVEY.md §8(c).  They guard against silent drift of the
oracle and are what the GPU parity tests compare against at these sizes.

    python tests/golden/make_golden.py
"""
import ctypes as C
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib  # noqa: E402

lib = oracle_lib.load()
out = {"source": "oracle/gpuprobe_oracle.c (parity unpinned: the reference holds no vectors)",
       "pattern": [], "digest": [], "generator": {}}
for nwords in (64, 4096, 64 * 1021, 1 << 20):
    for kind, passes in ((0, (0, 5, 63)), (1, (0, 31)), (2, (0, 1))):
        for pas in passes:
            r, _ = oracle_lib.pattern(lib, nwords, kind, pas)
            out["pattern"].append({"nwords": nwords, "kind": kind, "pass": pas,
                                   "result": [hex(v) for v in r.astuple()]})
for nwords in (4096, 1 << 18):
    passes = [(0, p) for p in range(64)] + [(2, 0), (2, 1)]
    d = 0
    for j, (kind, pas) in enumerate(passes):
        r, _ = oracle_lib.pattern(lib, nwords, kind, pas)
        d = lib.gp_cpu_digest_step(d, C.byref(r), j)
    out["digest"].append({"nwords": nwords, "passes": passes, "digest": hex(d)})
out["generator"] = {"seed": 0x5EED, "bf16": [int(x) for x in oracle_lib.gen_bf16(lib, 0x5EED, 0, 64)]}
with open(os.path.join(HERE, "pattern_golden.json"), "w") as f:
    json.dump(out, f, indent=1)
print("wrote", len(out["pattern"]), "pattern cases")
