N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N 2>gpurun_out/v12_${N}gpu.err | grep '^{' > gpurun_out/r1_v12_bench_cfg3_${N}gpu.json; python tools/bench_summary.py gpurun_out/r1_v12_bench_cfg3_${N}gpu.json
