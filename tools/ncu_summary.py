#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) into the few numbers DESIGN.md / profiles/ quote.  Usage: ncu_summary.py rep [kernel_idx]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units, data = rows[0], rows[1], rows[2:]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'launch__occupancy_limit_warps', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__cycles_active.avg', 'sm__cycles_elapsed.max']
d = data[idx]
print("kernel:", d[hdr.index('Kernel Name')])
for k in keys:
    if k in hdr: print(f"{k} = {d[hdr.index(k)]} {units[hdr.index(k)]}")
stall = [h for h in hdr if 'issue_stalled' in h and h.endswith('per_warp_active.pct')]
for v, h in sorted([(float(d[hdr.index(h)]), h) for h in stall], reverse=True)[:7]:
    print(f"{h} = {v}")
