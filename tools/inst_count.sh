#!/bin/bash
# usage: tools/inst_count.sh [workload]   — dynamic warp-instruction count and duration of the step kernel (ncu metrics only)
W=${1:-cfg3}
python tools/probe.py $W 3 > gpurun_out/ic_plain.log 2>&1 && ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none -k regex:step_ -s 6 -c 2 --csv --log-file gpurun_out/ic_$W.csv python tools/probe.py $W 3 > gpurun_out/ic_ncu.log 2>&1
grep -E "step_" gpurun_out/ic_$W.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | cut -c1-200
