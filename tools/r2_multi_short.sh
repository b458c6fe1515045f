#!/bin/bash
# usage (GPU box with N GPUs): tools/r2_multi_short.sh TAG N — the 2-rank peer-statistics test and the default bench line at N GPUs
# (weak scaling + the strong-scaling key) at the driver's --steps 20 and at the default step count
TAG=${1:-r2m}; N=${2:-8}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_peer_stats.py -q > gpurun_out/${TAG}_pytest_peer.log 2>&1; echo "pytest peer exit $?"; tail -2 gpurun_out/${TAG}_pytest_peer.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_cfg3_${N}gpu_steps20.json 2> gpurun_out/${TAG}_bench_cfg3_${N}gpu_steps20.err; echo "bench20 exit $?"
python tools/bench_summary.py gpurun_out/${TAG}_bench_cfg3_${N}gpu_steps20.json
timeout 600 $TR --master-port 29512 bench.py --gpus $N --no-cpu-baseline > gpurun_out/${TAG}_bench_cfg3_${N}gpu.json 2> gpurun_out/${TAG}_bench_cfg3_${N}gpu.err; echo "bench exit $?"
python tools/bench_summary.py gpurun_out/${TAG}_bench_cfg3_${N}gpu.json
