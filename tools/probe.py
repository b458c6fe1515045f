#!/usr/bin/env python
"""Timing probes for the step kernel on one GPU (CUDA events, 4 rotating env sets > L2): full step, step without obs,
obs only, move only.  Usage: probe.py [workload] [iters]"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "marl-drones_b200")); sys.path.insert(0, REPO)
import torch
from bench import WORKLOADS
from swarm_b200 import BatchedDroneEnv
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"; iters = int(sys.argv[2]) if len(sys.argv) > 2 else 40
n, kw = WORKLOADS[wl]; sets = 4
envs = [BatchedDroneEnv(n, auto_reset=True, seed=1, env_id_base=s * n, **kw) for s in range(sets)]
obs = [torch.empty(e.obs_shape, dtype=torch.float32, device="cuda") for e in envs]
acts = torch.randint(0, 4, (8, n, kw["n_drones"]), dtype=torch.uint8, device="cuda")
def timeit(fn, name):
    for i in range(5): fn(i)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i in range(iters):
        ev[i][0].record(); fn(i); ev[i][1].record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    print(f"{wl} {name:28s} median {ts[len(ts)//2]*1e3:8.1f} us   min {ts[0]*1e3:8.1f} us")
timeit(lambda i: envs[i % sets].step(acts[i % 8], out=obs[i % sets]), "step (all stages)")
timeit(lambda i: envs[i % sets].step(acts[i % 8], want_obs=False), "step, no obs")
timeit(lambda i: envs[i % sets].stage_obs(out=obs[i % sets]), "obs only (generic inst)")
timeit(lambda i: envs[i % sets].stage_move(acts[i % 8]), "move only")
timeit(lambda i: obs[i % sets].zero_(), "torch zero_ of obs")
timeit(lambda i: obs[i % sets].copy_(obs[(i + 1) % sets]), "torch copy_ obs->obs")
# the same step with uint8 observations (26 MB instead of 105 MB of obs per cfg3 step): how much of the step is obs bandwidth
envs8 = [BatchedDroneEnv(n, auto_reset=True, seed=1, env_id_base=s * n, obs_dtype=torch.uint8, **kw) for s in range(sets)]
obs8 = [torch.empty(e.obs_shape, dtype=torch.uint8, device="cuda") for e in envs8]
timeit(lambda i: envs8[i % sets].step(acts[i % 8], out=obs8[i % sets]), "step, uint8 obs")
# int32 actions (FAST instantiation, direct action loads instead of the bulk copy), and what dwell goals (generic instantiation) cost
acts32 = acts.to(torch.int32)
timeit(lambda i: envs[i % sets].step(acts32[i % 8], out=obs[i % sets]), "step, int32 actions")
envsd = [BatchedDroneEnv(n, auto_reset=True, seed=1, env_id_base=s * n, dwell_normal=10, dwell_anom=35, **kw) for s in range(sets)]
timeit(lambda i: envsd[i % sets].step(acts[i % 8], out=obs[i % sets]), "step, dwell 10/35")
timeit(lambda i: envsd[i % sets].step(acts32[i % 8], out=obs[i % sets]), "step, dwell, int32 actions")
