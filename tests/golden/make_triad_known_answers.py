#!/usr/bin/env python3
"""Known answers for the triad probe (SUM / XOR of the u64 words of a = fmaf(3, c, b), seed 0x5EED) computed by the
CPU oracle's streaming gp_cpu_triad_checksum, for the array sizes the plugin is likely to run with.
    python tests/golden/make_triad_known_answers.py > tests/golden/triad_known_answers.json
(the 16 GiB entry takes a few seconds on 8 cores)"""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import oracle_lib
lib = oracle_lib.load()
out = {"seed": "0x5EED", "scalar": 3.0, "what": "sum / xor (mod 2^64) of the u64 words of the triad result, by bytes per array",
       "made_by": "tests/golden/make_triad_known_answers.py (oracle/gpuprobe_oracle.c gp_cpu_triad_checksum)", "answers": {}}
for nbytes in (64 << 20, 256 << 20, 1 << 30, 4 << 30, 16 << 30):
    s, x = oracle_lib.triad_checksum(lib, 0x5EED, nbytes // 4, 3.0, nthreads=os.cpu_count() or 1)
    out["answers"][str(nbytes)] = {"sum": hex(s), "xor": hex(x)}
print(json.dumps(out, indent=1))
