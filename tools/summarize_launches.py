#!/usr/bin/env python3
"""ncu launch list (gpu__time_duration.sum CSV) -> per-kernel share of the bench step, as markdown.
   python tools/summarize_launches.py gpurun_out/launches_r02.csv 5 > profiles/launches_r02_summary.md"""
import collections, csv, sys
path, steps = sys.argv[1], int(sys.argv[2])
rows = [r for r in csv.reader(open(path)) if len(r) > 5]
hdr = rows[0]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    name = r[ik].split("(")[0].replace("void ", "")
    v = float(r[iv].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6}[r[iu]]
    agg.setdefault(name, []).append(v)
step = {k: v for k, v in agg.items() if len(v) % steps == 0 and not k.startswith(("gen_", "verify_scratch"))}
tot = sum(sum(v) for v in step.values())
print("| kernel | launches/step | ns/step | share of step |\n|---|---|---|---|")
for k, v in step.items():
    print(f"| `{k}` | {len(v) // steps} | {sum(v) / steps:.0f} | {sum(v) / tot:.3f} |")
print(f"| total | {sum(len(v) for v in step.values()) // steps} | {tot / steps:.0f} | 1.000 |")
other = {k: v for k, v in agg.items() if k not in step}
if other:
    print("\nOutside the step (set-up, once): " + ", ".join(f"`{k}` x{len(v)} ({sum(v) / 1e3:.0f} us)" for k, v in other.items()))
