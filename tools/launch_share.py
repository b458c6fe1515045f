#!/usr/bin/env python
"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) by kernel: launches, total time and share.
usage: launch_share.py launches.csv [last_n]   — last_n: only the last n launches (e.g. the steady-state tail of a run)."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr, rows = rows[0], rows[1:]
kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
if len(sys.argv) > 2:
    rows = rows[-int(sys.argv[2]):]
tot = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name = re.sub(r"\(.*", "", r[kn])
    name = re.sub(r"<.*", "", name) if not name.startswith("swarm::") else name
    tot[name][0] += 1
    tot[name][1] += float(r[mv].replace(",", ""))
all_ns = sum(v[1] for v in tot.values())
print(f"{len(rows)} launches, {all_ns / 1e6:.3f} ms")
print("| kernel | launches | ms | share |\n|---|---|---|---|")
for name, (n, ns) in sorted(tot.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"| `{name}` | {n} | {ns / 1e6:.3f} | {ns / all_ns:.3f} |")
