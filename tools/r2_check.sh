#!/bin/bash
# usage (GPU box): tools/r2_check.sh TAG [full] — GPU tests, bench at the driver's --steps 20 and at the default, ncu launch list;
# "full" adds one ncu --set full capture (with source) of the cfg3 step kernel.  Everything lands under gpurun_out/.
TAG=${1:-r2}; FULL=${2:-}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_cfg3_steps20.json 2> gpurun_out/${TAG}_bench_cfg3_steps20.err; echo "bench20 exit $?"
python tools/bench_summary.py gpurun_out/${TAG}_bench_cfg3_steps20.json
python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench_cfg3.json 2> gpurun_out/${TAG}_bench_cfg3.err; echo "bench exit $?"
python tools/bench_summary.py gpurun_out/${TAG}_bench_cfg3.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_cfg3.csv \
    python bench.py --steps 64 --warmup 3 --no-graph --no-cpu-baseline --no-strong --clock-window-s 0 --e2e-steps 2 > gpurun_out/${TAG}_ncu_launch.log 2>&1
echo "ncu launches exit $?"
if [ -n "$FULL" ]; then
  ncu --set full --clock-control none --import-source on -k regex:step_warp -s 6 -c 1 -o gpurun_out/${TAG}_cfg3 -f python tools/probe.py cfg3 3 > gpurun_out/${TAG}_ncu_full.log 2>&1
  echo "ncu full exit $?"
fi
