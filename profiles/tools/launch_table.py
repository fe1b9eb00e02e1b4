#!/usr/bin/env python
"""Print the per-launch table (time, warp instructions, kernel) of an ncu --csv launch list:
python profiles/tools/launch_table.py gpurun_out/<name>_launches.csv"""
import collections
import csv
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]
ki, mi, vi, ii = h.index('Kernel Name'), h.index('Metric Name'), h.index('Metric Value'), h.index('ID')
d = collections.OrderedDict()
for r in rows[1:]:
    d.setdefault(r[ii], {'k': r[ki]})[r[mi]] = float(r[vi].replace(',', ''))
tot = 0
for v in d.values():
    t = v['gpu__time_duration.sum'] / 1000
    tot += t
    print(f"{t:8.1f} us {v.get('smsp__inst_executed.sum', 0) / 1e6:7.2f} Minst  {v['k'][:90]}")
print(f"{tot:8.1f} us total")
