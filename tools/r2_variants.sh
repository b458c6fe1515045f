#!/bin/bash
# usage (GPU box): tools/r2_variants.sh TAG — GPU tests (no -x), then the cfg3 bench under the staged kernel's tuning knobs
TAG=${1:-r2v}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --maxfail=25 -x --deselect tests/test_gpu_unchanged_callers.py > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -30 gpurun_out/${TAG}_pytest.log
B="python bench.py --no-cpu-baseline --no-strong --e2e-steps 2 --clock-window-s 0.1"
for v in "SWARM_DYN=0" "SWARM_DYN=1" "SWARM_DYN=0 SWARM_WPB=4" "SWARM_DYN=1 SWARM_WPB=4" "SWARM_DYN=0 SWARM_STAGES=2" "SWARM_DYN=1 SWARM_STAGES=2"; do
  n=$(echo $v | tr ' =' '__')
  env $v $B > gpurun_out/${TAG}_bench_${n}.json 2> gpurun_out/${TAG}_bench_${n}.err; echo "$v exit $?"
  python tools/bench_summary.py gpurun_out/${TAG}_bench_${n}.json
done
