#!/bin/bash
# usage (GPU box): tools/prof.sh lib.so TAG [workload] — ncu --set full capture (with source) of the step kernel of a library variant
so=$1; tag=$2; wl=${3:-cfg3}
export SWARM_B200_SO=$PWD/$so
mkdir -p gpurun_out
python tools/probe.py $wl 20 > gpurun_out/${tag}_probe.log 2>&1; head -3 gpurun_out/${tag}_probe.log
ncu --set full --clock-control none --import-source on -k regex:step_ -s 6 -c 1 -o gpurun_out/${tag} -f python tools/probe.py $wl 3 > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu exit $?"
