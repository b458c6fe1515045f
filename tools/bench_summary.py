#!/usr/bin/env python
"""One line per bench JSON line found in the given files."""
import json, sys
for f in sys.argv[1:]:
    for l in open(f):
        if l.startswith("{"):
            d = json.loads(l); r = d.get("roofline", {})
            print(f"{f}: {d['config'].get('workload')} value={d['value']:.3e} kern_ms={r.get('kernel_ms_avg', 0):.4f} frac={r.get('frac', 0):.3f} e2e={d['e2e']['value']:.3e}")
