#!/bin/bash
# usage (GPU box): tools/sweep.sh [ENVVAR=val ...] — GPU parity tests, the cfg3 bench line with the default settings, then once per given env setting
mkdir -p gpurun_out
B="python bench.py --no-cpu-baseline --e2e-steps 2 --clock-window-s 0.1"
python -m pytest tests/test_gpu_parity.py tests/test_gpu_surface.py -x -q 2>&1 | tail -2
$B > gpurun_out/sw_default.json 2> gpurun_out/sw_default.err; echo "default: $(python tools/bench_summary.py gpurun_out/sw_default.json)"
for cfg in "$@"; do
  tag=$(echo $cfg | tr '= ' '__')
  env $cfg $B > gpurun_out/sw_$tag.json 2> gpurun_out/sw_$tag.err; echo "$cfg: $(python tools/bench_summary.py gpurun_out/sw_$tag.json)"
done
