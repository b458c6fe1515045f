#!/usr/bin/env python
"""Hot-spot view of `ncu --page source --print-source sass --csv`: runs of SASS with equal execution counts.
Usage: sass_hot.py file.csv [min_pct]"""
import csv, itertools, sys
rows = list(csv.reader(open(sys.argv[1]))); minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
his = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hdr = rows[his[0]]; end = his[1] - 1 if len(his) > 1 else len(rows)
data = [r for r in rows[his[0] + 1:end] if len(r) == len(hdr)]
ie, src, smp = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
tot = sum(int(r[ie]) for r in data); ts = sum(int(r[smp]) for r in data)
print('total warp inst', tot, 'sass rows', len(data), 'samples', ts)
for k, g in itertools.groupby(data, key=lambda r: int(r[ie])):
    g = list(g); n = len(g); sm = sum(int(r[smp]) for r in g)
    if k * n > tot * minp / 100 or sm > ts * minp / 100:
        print(f"{k:>9} x{n:>4} = {k*n/tot*100:5.1f}% inst {sm/ts*100:5.1f}% samples  @{g[0][0][-5:]} {g[0][src][:60]}")
