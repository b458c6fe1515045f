#!/bin/bash
# usage (GPU box): tools/r2_variants.sh TAG "VAR=.. VAR=.." "VAR=.." ... — GPU tests (without the slow unchanged-caller runs), then the
# cfg3 bench under each given set of environment knobs (SWARM_WPB / SWARM_STAGES / SWARM_DYN / SWARM_B200_SO=variant library,
# BENCH_ARGS="extra bench.py flags")
TAG=${1:-r2v}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_unchanged_callers.py > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -30 gpurun_out/${TAG}_pytest.log
for v in "$@"; do
  n=$(echo $v | sed 's#SWARM_B200_SO=[^ ]*/libswarm_\([A-Za-z0-9]*\).so#\1#; s#BENCH_ARGS=##' | tr ' =-' '___')
  extra=$(echo "$v" | grep -o 'BENCH_ARGS=.*' | sed 's/BENCH_ARGS=//')
  vars=$(echo "$v" | sed 's/BENCH_ARGS=.*//')
  env $vars python bench.py --no-cpu-baseline --no-strong --e2e-steps 2 --clock-window-s 0.1 $extra > gpurun_out/${TAG}_bench_${n}.json 2> gpurun_out/${TAG}_bench_${n}.err; echo "$v exit $?"
  python tools/bench_summary.py gpurun_out/${TAG}_bench_${n}.json
done
