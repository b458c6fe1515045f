#!/bin/bash
# usage (GPU box with N GPUs): tools/r2_multi.sh TAG N — the 2-rank peer-statistics test, the default bench line at N GPUs (weak
# scaling + the strong-scaling key), the NCCL statistics form, and the host-copy ceiling of the box
TAG=${1:-r2m}; N=${2:-2}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_peer_stats.py -q > gpurun_out/${TAG}_pytest_peer.log 2>&1; echo "pytest peer exit $?"; tail -3 gpurun_out/${TAG}_pytest_peer.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_cfg3_${N}gpu_steps20.json 2> gpurun_out/${TAG}_bench_cfg3_${N}gpu_steps20.err; echo "bench20 exit $?"
python tools/bench_summary.py gpurun_out/${TAG}_bench_cfg3_${N}gpu_steps20.json
timeout 600 $TR --master-port 29512 bench.py --gpus $N > gpurun_out/${TAG}_bench_cfg3_${N}gpu.json 2> gpurun_out/${TAG}_bench_cfg3_${N}gpu.err; echo "bench exit $?"
python tools/bench_summary.py gpurun_out/${TAG}_bench_cfg3_${N}gpu.json
timeout 600 $TR --master-port 29513 bench.py --gpus $N --stats-mode nccl --no-strong > gpurun_out/${TAG}_bench_cfg3_${N}gpu_nccl.json 2> gpurun_out/${TAG}_bench_cfg3_${N}gpu_nccl.err; echo "bench nccl exit $?"
python tools/bench_summary.py gpurun_out/${TAG}_bench_cfg3_${N}gpu_nccl.json
timeout 600 $TR --master-port 29514 bench.py --gpus $N --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_ref_${N}gpu.json 2> gpurun_out/${TAG}_ref_${N}gpu.err; echo "reference arm exit $?"
timeout 300 $TR --master-port 29515 tools/pcie_ceiling.py > gpurun_out/${TAG}_pcie_${N}gpu.json 2> gpurun_out/${TAG}_pcie_${N}gpu.err; echo "pcie exit $?"; cat gpurun_out/${TAG}_pcie_${N}gpu.json
