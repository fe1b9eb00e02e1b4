#!/usr/bin/env python
"""Digest of `ncu -i x.ncu-rep --page raw --csv` (+ optional `--page source --csv --print-source sass`):
headline metrics, top stall reasons per kernel, and the SASS lines with the most stall samples.
python profiles/tools/ncu_digest.py raw.csv [source.csv]"""
import csv
import sys


def f(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return 0.0


rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
want = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__waves_per_multiprocessor', 'launch__grid_size',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__cycles_active.avg']
for r in rows[2:]:
    print(r[h.index('Kernel Name')][:100])
    for w in want:
        if w in h:
            print(f"  {w:75s} {r[h.index(w)]} {rows[1][h.index(w)]}")
    st = [(h[i], r[i]) for i in range(len(h)) if 'issue_stalled' in h[i] and h[i].endswith('per_issue_active.ratio')]
    for k, v in sorted(st, key=lambda kv: -f(kv[1]))[:7]:
        print('     stall', k.replace('smsp__average_warps_issue_stalled_', '').replace('smsp__average_warp_latency_issue_stalled_', ''), v)
if len(sys.argv) > 2:
    rows = list(csv.reader(open(sys.argv[2])))
    hdr = [i for i, r in enumerate(rows) if 'Source' in r and '# Samples' in r] + [len(rows)]
    for a, b in zip(hdr[:-1], hdr[1:]):
        h = rows[a]
        si, j = h.index('Source'), h.index('Warp Stall Sampling (All Samples)')
        body = [r for r in rows[a + 1:b] if len(r) > j]
        tot = sum(f(r[j]) for r in body) or 1
        print(f'--- kernel at row {a}: {tot:.0f} samples, {len(body)} SASS lines')
        top = sorted([(f(r[j]), k, r[si]) for k, r in enumerate(body)], reverse=True)[:int(sys.argv[3]) if len(sys.argv) > 3 else 14]
        for s, k, src in sorted(top, key=lambda x: x[1]):
            print(f"{s:7.0f} {100 * s / tot:5.1f}%  {k:5d}  {src[:110]}")
