#!/usr/bin/env python
"""Turn `ncu -i X.ncu-rep --page raw --csv` output into the per-kernel table kept under profiles/.
usage: python profiles/summarize_ncu.py raw.csv "title" > profiles/rNN_callK_full.md"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
cols = [("us", "gpu__time_duration.sum"), ("regs", "launch__registers_per_thread"),
        ("warps %", "sm__warps_active.avg.pct_of_peak_sustained_active"),
        ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        ("fmaheavy %", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"),
        ("alu %", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
        ("dram rd MB", "dram__bytes_read.sum"), ("dram wr MB", "dram__bytes_write.sum"),
        ("dram %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        ("Minst", "smsp__inst_executed.sum")]
units = rows[1]
print(f"# {sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]}\n")
print("| kernel | " + " | ".join(c for c, _ in cols) + " |")
print("|---|" + "---|" * len(cols))


def val(r, m):
    v = float(r[ix[m]])
    u = units[ix[m]].lower()
    if m.startswith("dram__bytes"):
        v *= {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1.0)
    if m == "smsp__inst_executed.sum":
        v /= 1e6
    if m == "gpu__time_duration.sum" and u.startswith("ns"):
        v /= 1e3
    if m == "gpu__time_duration.sum" and u.startswith("ms"):
        v *= 1e3
    return v


tot = 0.0
for r in data:
    name = r[ix["Kernel Name"]].replace("void ", "").replace("<unnamed>::", "")
    name = name.split("(")[0]
    v = [val(r, m) for _, m in cols]
    tot += v[0]
    print(f"| {name} | " + " | ".join(f"{x:.1f}" for x in v) + " |")
print(f"\nsum of kernel times (cold-cache, serialised under ncu): {tot:.1f} us")
