#!/usr/bin/env python
"""Per-launch listing of the last full train step from an `ncu --metrics gpu__time_duration.sum --csv` log."""
import collections, csv, io, re, sys
path = sys.argv[1]
lines = open(path).read().splitlines()
i = next(k for k, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[i:]))))
names = [re.sub(r'^(void )?<unnamed>::', '', r['Kernel Name']).split('(')[0] for r in rows]
# a step runs from one weight re-layout launch (first entry of the schedule) to the next; the step has two adam_k launches
ad = [k for k, n in enumerate(names) if n.startswith('weight_prep')]
s, e = ad[-2], ad[-1]
tot = 0
agg = collections.defaultdict(lambda: [0, 0.0])
out = []
for k in range(s, e):
    r = rows[k]
    v = float(r['Metric Value'].replace(',', ''))
    us = v / 1000 if r['Metric Unit'] in ('ns', 'nsecond') else v
    tot += us; agg[names[k]][0] += 1; agg[names[k]][1] += us
    out.append(f"{k-s:3d} {names[k][:30]:30s} grid={r['Grid Size']:>14s} blk={r['Block Size']:>12s} {us:7.1f}")
if "-v" in sys.argv:
    print("\n".join(out))
print(f"{e-s} kernels, {tot:.1f} us summed")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n[:44]:44s} {c:3d} {t:8.1f} {100*t/tot:5.1f}%")
