#!/bin/bash
# usage (GPU box): tools/sweep_libs.sh lib1.so ... — the cfg3 bench line for each kernel-variant library (SWARM_B200_SO override)
mkdir -p gpurun_out
B="python bench.py --no-cpu-baseline --e2e-steps 2 --clock-window-s 0.1"
for so in "$@"; do
  tag=$(basename $so .so); SWARM_B200_SO=$PWD/$so $B > gpurun_out/sw_$tag.json 2> gpurun_out/sw_$tag.err; echo "$tag: $(python tools/bench_summary.py gpurun_out/sw_$tag.json)"
done
