#!/usr/bin/env python
"""Per-source-line instruction / stall-sample attribution from `ncu --page source --print-source cuda,sass --csv`.
Usage: src_hot.py file.csv [topN]   (first profiled kernel only)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]))); top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = []; cur = None; hdr = None; fn = None; first_fn = None
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        fn = r[1]; first_fn = first_fn or fn; continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr and fn == first_fn and r[0] not in ('', '-') and len(r) >= 9:
        try: out.append((int(r[7]), int(r[6]), cur, r[0], r[1].strip()[:100]))
        except ValueError: pass
# the same kernel may be listed for 2 captures: dedupe identical (file,line) keeping the first
seen = {}; 
for o in out: seen.setdefault((o[2], o[3]), o)
out = list(seen.values())
tot = sum(o[0] for o in out); ts = sum(o[1] for o in out)
print("warp inst", tot, "samples", ts)
for o in sorted(out, reverse=True)[:top]:
    print(f"{o[0]/tot*100:5.1f}% inst {o[1]/max(ts,1)*100:5.1f}% smp  {o[2]}:{o[3]}  {o[4]}")
