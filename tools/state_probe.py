#!/usr/bin/env python
"""Time the non-step entry points (reset, state extraction / injection) next to their byte counts.  Usage: state_probe.py [workload ...]"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "marl-drones_b200")); sys.path.insert(0, REPO)
import torch
from bench import WORKLOADS
from swarm_b200 import BatchedDroneEnv

def t_us(fn, n=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3

for wl in (sys.argv[1:] or ["cfg4", "cfg2", "cfg3", "cfg5a"]):
    n, kw = WORKLOADS[wl]
    env = BatchedDroneEnv(n, auto_reset=True, seed=1, **kw)
    a = torch.randint(0, 4, (n, kw["n_drones"]), dtype=torch.uint8, device="cuda")
    for _ in range(50): env.step(a, want_obs=False)
    g, p, t, w = env.get_state()
    mb = g.numel() / 1e6
    res = dict(step=t_us(lambda: env.step(a)), reset_all=t_us(lambda: env.reset()), get_state=t_us(lambda: env.get_state()),
               set_state=t_us(lambda: env.set_state(g, p, t, w)))
    print(wl, f"int8 grid {mb:.1f} MB |", " | ".join(f"{k} {v:8.1f} us" for k, v in res.items()), flush=True)
    env.close()
