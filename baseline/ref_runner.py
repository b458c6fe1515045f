"""Times the UNMODIFIED reference `DroneEnv` (baseline/_ref/A2C/noVis_sim.py — a verbatim copy of
/root/reference/A2C/noVis_sim.py installed by `__graft_entry__.build()`, git-ignored, shipped to the GPU box like the
built .so) on the host cores: P worker processes, each stepping its own envs with random actions and reset-on-done, no
per-env-step IPC (SURVEY §8(d) "Primary" CPU baseline; the loop body is A2C/envs.py:22-26).

Nothing of this repo's engine, kernels or oracle is on this path: the workers import only numpy and the reference module.

    python baseline/ref_runner.py --steps 3 --warmup 1 --row_count 64 --col_count 64 --n_drones 16 --step_size 1.0

prints one JSON object: {"value": agent-steps/s, "procs": P, "envs_per_proc": E, "inner": I, "steps": K, "wall_s": ...}.
A "step" = every worker advances each of its E envs by I env-steps (a bounded sample of the workload shape).
"""
import argparse
import importlib.util
import json
import multiprocessing as mp
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
VARIANTS = {"A2C": os.path.join("A2C", "noVis_sim.py"), "DQL": os.path.join("DQL_model", "noVis_sim.py")}


def available(variant="A2C"):
    return os.path.exists(os.path.join(REF_DIR, VARIANTS[variant]))


def load_reference(variant="A2C"):
    """The reference module, imported from the installed copy under baseline/_ref (never from this repo's package)."""
    path = os.path.join(REF_DIR, VARIANTS[variant])
    spec = importlib.util.spec_from_file_location(f"_installed_ref_noVis_sim_{variant}", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _worker(conn, wid, variant, kw, envs_per_proc):
    import numpy as np
    mod = load_reference(variant)
    np.random.seed(10_000 + wid)                       # the reference draws from the global legacy np.random
    envs = [mod.DroneEnv(**kw) for _ in range(envs_per_proc)]
    for e in envs:
        e.reset()
    rs = np.random.RandomState(1234 + wid)
    D = kw["n_drones"]
    conn.send("ready")
    while True:
        msg = conn.recv()
        if msg is None:
            break
        inner = msg
        t0 = time.perf_counter()
        for _ in range(inner):
            for e in envs:
                ob, reward, done = e.step(list(rs.randint(0, 4, D)))
                if done:                               # A2C/envs.py:24-25
                    ob = e.reset()
        conn.send(time.perf_counter() - t0)
    conn.close()


class ReferencePool:
    """P processes holding reference envs; `step(inner)` advances all of them and returns the wall time of the slowest."""

    def __init__(self, kw, procs, envs_per_proc=1, variant="A2C"):
        ctx = mp.get_context("fork" if "torch" not in sys.modules else "spawn")
        self.kw, self.procs, self.envs_per_proc = dict(kw), procs, envs_per_proc
        self.conns, self.ps = [], []
        for w in range(procs):
            a, b = ctx.Pipe()
            p = ctx.Process(target=_worker, args=(b, w, variant, self.kw, envs_per_proc), daemon=True)
            p.start()
            self.conns.append(a)
            self.ps.append(p)
        for c in self.conns:
            assert c.recv() == "ready"

    def step(self, inner):
        t0 = time.perf_counter()
        for c in self.conns:
            c.send(inner)
        for c in self.conns:
            c.recv()
        return time.perf_counter() - t0

    def close(self):
        for c in self.conns:
            try:
                c.send(None)
            except Exception:
                pass
        for p in self.ps:
            p.join(timeout=5)
            if p.is_alive():
                p.terminate()


def host_cores():
    """Cores this process may run on (ignores OMP_NUM_THREADS, which torchrun sets to 1)."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def measure(kw, steps, warmup, budget_s=45.0, procs=0, envs_per_proc=1, variant="A2C", inner=0):
    """Returns a dict with agent-steps/s of the unmodified reference on `procs` processes (default: every host core)."""
    procs = procs or host_cores()
    pool = ReferencePool(kw, procs, envs_per_proc, variant)
    try:
        dt = pool.step(1)                               # calibration: one env-step of every env (also warms the workers)
        if not inner:
            inner = max(1, min(100000, int(budget_s / max(steps + warmup, 1) / max(dt, 1e-6))))
        for _ in range(warmup):
            pool.step(inner)
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.step(inner)
        wall = time.perf_counter() - t0
    finally:
        pool.close()
    agent_steps = float(procs) * envs_per_proc * inner * steps * kw["n_drones"]
    return {"value": agent_steps / wall, "procs": procs, "envs_per_proc": envs_per_proc, "inner": inner, "steps": steps,
            "warmup": warmup, "wall_s": wall, "variant": variant,
            "source": "baseline/_ref/" + VARIANTS[variant].replace(os.sep, "/") + " (unmodified copy of the reference)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--budget-s", type=float, default=45.0)
    ap.add_argument("--procs", type=int, default=0)
    ap.add_argument("--envs-per-proc", type=int, default=1)
    ap.add_argument("--inner", type=int, default=0)
    ap.add_argument("--variant", default="A2C", choices=sorted(VARIANTS))
    ap.add_argument("--row_count", type=int, default=10)
    ap.add_argument("--col_count", type=int, default=10)
    ap.add_argument("--n_drones", type=int, default=3)
    ap.add_argument("--n_anamolous", type=int, default=5)
    ap.add_argument("--step_size", type=float, default=0.5)
    a = ap.parse_args()
    if not available(a.variant):
        print(json.dumps({"unavailable": "baseline/_ref is not installed (run __graft_entry__.build() where /root/reference exists)"}))
        return
    kw = dict(row_count=a.row_count, col_count=a.col_count, n_drones=a.n_drones, n_anamolous=a.n_anamolous, step_size=a.step_size)
    print(json.dumps(measure(kw, a.steps, a.warmup, a.budget_s, a.procs, a.envs_per_proc, a.variant, a.inner)))


if __name__ == "__main__":
    main()
