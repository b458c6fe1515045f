#!/usr/bin/env python
"""Per-step latency of the single-env facade (what the reference's unchanged callers see) next to the unmodified reference
DroneEnv from baseline/_ref on a host core.  Usage: facade_latency.py [steps]"""
import importlib.util, json, os, sys, time
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "marl-drones_b200")); sys.path.insert(0, REPO)
import numpy as np
from swarm_b200 import DroneEnv
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
ref_path = os.path.join(REPO, "baseline", "_ref", "A2C", "noVis_sim.py")
RefEnv = None
if os.path.exists(ref_path):
    spec = importlib.util.spec_from_file_location("ref_noVis_sim", ref_path); m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    RefEnv = m.DroneEnv
out = {}
for name, kw in (("trainer_5x5_d2", dict(row_count=5, col_count=5, n_drones=2, step_size=1.0, n_anamolous=2)),
                 ("defaults_10x10_d3", dict()),
                 ("cfg4_shape_10x10_d8", dict(row_count=10, col_count=10, n_drones=8, step_size=1.0)),
                 ("cfg2_shape_32x32_d4", dict(row_count=32, col_count=32, n_drones=4, step_size=1.0)),
                 ("cfg3_shape_64x64_d16", dict(row_count=64, col_count=64, n_drones=16, step_size=1.0))):
    res = {}
    for label, cls in (("facade_gpu", DroneEnv), ("reference_cpu", RefEnv)):
        if cls is None:
            continue
        np.random.seed(1)
        env = cls(**kw)
        D = env.n_drones
        rs = np.random.RandomState(0)
        acts = rs.randint(0, 4, (steps, D))
        n = steps if label == "facade_gpu" else max(50, min(steps, int(4e6 / (env.state_size * D))))
        for k in range(20):
            env.step(list(acts[k]))
        t0 = time.perf_counter()
        for k in range(n):
            _, _, done = env.step(list(acts[k]))
            if done:
                env.reset()
        res[label + "_us_per_step"] = round((time.perf_counter() - t0) / n * 1e6, 1)
    out[name] = res
    print(name, res, flush=True)
print(json.dumps(out))
