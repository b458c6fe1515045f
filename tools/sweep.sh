#!/bin/bash
# usage (GPU box): tools/sweep.sh — parity tests, then the cfg3 bench line (optionally for several libraries given as arguments)
mkdir -p gpurun_out
B="python bench.py --no-cpu-baseline --e2e-steps 2 --clock-window-s 0.1"
python -m pytest tests/test_gpu_parity.py tests/test_gpu_surface.py -x -q 2>&1 | tail -2
$B > gpurun_out/sw_default.json 2> gpurun_out/sw_default.err; python tools/bench_summary.py gpurun_out/sw_default.json
for so in "$@"; do
  tag=$(basename $so .so); SWARM_B200_SO=$PWD/$so $B > gpurun_out/sw_$tag.json 2> gpurun_out/sw_$tag.err; python tools/bench_summary.py gpurun_out/sw_$tag.json
done
