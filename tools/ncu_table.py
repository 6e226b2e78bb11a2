#!/usr/bin/env python3
"""ncu --page raw --csv -> one markdown row per captured launch with the metrics DESIGN.md quotes.
   ncu -i prof.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_table.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
cols = [("duration", "gpu__time_duration.sum"), ("dram read", "dram__bytes_read.sum"), ("dram write", "dram__bytes_write.sum"),
        ("dram active %", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed"),
        ("L2 % ", "lts__throughput.avg.pct_of_peak_sustained_elapsed"), ("L2 hit %", "lts__t_sector_hit_rate.pct"),
        ("SM %", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("issue active %", "sm__inst_issued.avg.pct_of_peak_sustained_active"),
        ("warps active %", "sm__warps_active.avg.pct_of_peak_sustained_active"),
        ("tensor pipe active %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
        ("regs", "launch__registers_per_thread"), ("grid", "launch__grid_size"), ("block", "launch__block_size"),
        ("SM clock", "sm__cycles_elapsed.avg.per_second")]
ik = hdr.index("Kernel Name")
print("| kernel | " + " | ".join(c[0] for c in cols) + " |")
print("|---|" + "---|" * len(cols))
for r in rows[2:]:
    name = r[ik].split("(")[0].replace("void ", "").replace("gp::", "")
    cells = []
    for _, m in cols:
        if m not in hdr:  # this ncu prefixes some section metrics (e.g. FBSP.TriageCompute.dram__throughput...)
            alt = [h for h in hdr if h.endswith("." + m)]
            m = alt[0] if alt else m
        if m in hdr:
            i = hdr.index(m)
            try:
                v = float(r[i].replace(",", ""))
                cells.append(f"{v:.4g} {units[i]}".strip())
            except ValueError:
                cells.append(r[i])
        else:
            cells.append("-")
    print(f"| `{name}` | " + " | ".join(cells) + " |")
