#!/usr/bin/env python
"""Time swarm_greedy_actions (the scripted nearest-unmapped-cell action source) next to the env step it feeds, early and late in the
episodes.  Usage: greedy_probe.py [workload ...]"""
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

for wl in (sys.argv[1:] or ["cfg4", "cfg2", "cfg3", "cfg5b"]):
    n, kw = WORKLOADS[wl]
    env = BatchedDroneEnv(n, auto_reset=False, seed=1, **kw)
    out = []
    for phase, steps in (("fresh", 0), ("after 64 greedy steps", 64), ("after 256", 192), ("after 1024", 768)):
        for _ in range(steps):
            env.step(env.greedy_actions(), want_obs=False)
        left = float(env.meta[:, 1].float().mean()) / (kw["row_count"] * kw["col_count"])
        out.append(f"{phase}: greedy {t_us(env.greedy_actions):9.1f} us  (unmapped {left:.2f})")
    a = env.greedy_actions()
    print(wl, "| step", f"{t_us(lambda: env.step(a), 20):7.1f} us |", " | ".join(out), flush=True)
    env.close()
