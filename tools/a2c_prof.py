#!/usr/bin/env python
"""Where the batched A2C iteration (bench.py --workload cfg4_a2c) spends its device time: torch.profiler kernel table.
Usage (GPU box): python tools/a2c_prof.py [--tf32]"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "marl-drones_b200")); sys.path.insert(0, REPO)
import torch
from torch.profiler import profile, ProfilerActivity
from bench import WORKLOADS
from swarm_b200 import BatchedDroneEnv
from swarm_b200.a2c import A2CTrainer
if "--tf32" in sys.argv:
    torch.backends.cuda.matmul.allow_tf32 = True; torch.backends.cudnn.allow_tf32 = True
n, kw = WORKLOADS["cfg4"]
env = BatchedDroneEnv(n, auto_reset=True, seed=2026, **kw)
tr = A2CTrainer(env, rollout_steps=25, seed=99)
for _ in range(2): tr.train_iteration()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record(); tr.train_iteration(); ev[1].record(); torch.cuda.synchronize()
print("iteration: %.1f ms" % ev[0].elapsed_time(ev[1]))
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    tr.train_iteration(); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=70))
