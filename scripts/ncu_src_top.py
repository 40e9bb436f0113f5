#!/usr/bin/env python
"""Top stall sites + code-region buckets from `ncu --page source --csv` (and headline metrics from `--page raw --csv`)."""
import csv, sys
from collections import Counter
src, raw = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(src))); hdr = rows[1]; data = rows[2:]
iS, iI, iSrc = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Source')
tot = sum(int(r[iS]) for r in data)
print('samples', tot, 'warp instr', sum(int(r[iI]) for r in data), 'sass lines', len(data))
B = 100
for b in range(0, len(data), B):
    ss = sum(int(r[iS]) for r in data[b:b + B]); ii = sum(int(r[iI]) for r in data[b:b + B])
    key = [r[iSrc].strip().split()[0] if not r[iSrc].strip().startswith('@') else r[iSrc].strip().split()[1] for r in data[b:b + B]]
    marks = [k for k in key if k.startswith(('UTC', 'UTMA', 'LDTM', 'SYNCS.PHASE', 'LDS', 'STS', 'STG', 'LDG', 'BAR', 'F2FP', 'RED', 'ATOM', 'FFMA2', 'MUFU'))]
    if ss > tot * 0.01:
        print(b, ss, f"{100 * ss / tot:.1f}%", ii, dict(Counter(marks).most_common(5)))
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
for k in sorted(range(len(data)), key=lambda k: -int(data[k][iS]))[:14]:
    r = data[k]; st = sorted([(int(r[i]), hdr[i]) for i in stall_cols], reverse=True)[:2]
    print(k, r[iS], r[iI], r[iSrc].strip()[:64], st)
r = list(csv.reader(open(raw))); h = r[0]; v = r[2]
for w in ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size', 'smsp__inst_executed.sum']:
    if w in h: print(w, v[h.index(w)], r[1][h.index(w)])
