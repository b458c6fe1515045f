"""Times the UNMODIFIED reference training loops on the host cores — the reference arm of bench.py's caller-row workloads:

  a2c  A2C/train.py::train() with A2C/models.Policy (MLP) nets on A2C/noVis_sim.DroneEnv (10x10, 8 drones): a "step" = one outer
       iteration (rollout_steps env steps with the policy in the loop, returns, loss, backward, clipped AdamW update);
  dql  DQL_model/DQLearning.py: Agent.act -> env.step -> memorize -> replay(32), driven exactly as its module-level loop does
       (`:161-187`); a "step" = one pass of that loop.

Everything is imported from baseline/_ref (verbatim copies installed by __graft_entry__.build()); nothing of this repo's
engine is on the path.  The reference is single-process, so "all host cores" = P independent copies of the loop, one per core
(torch intra-op threads = 1 each).  Prints one JSON object.

    python baseline/ref_trainers.py a2c --steps 3 --warmup 1
"""
import argparse
import json
import logging
import multiprocessing as mp
import os
import sys
import tempfile
import time
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def available():
    return os.path.exists(os.path.join(REF, "A2C", "train.py")) and os.path.exists(os.path.join(REF, "DQL_model", "DQLearning.py"))


def _a2c_worker(conn, wid, grid, n_drones, T, warmup, steps):
    import numpy as np
    import torch
    torch.set_num_threads(1)
    sys.path.insert(0, os.path.join(REF, "A2C"))
    os.chdir(tempfile.mkdtemp())                       # train() writes its log file into the cwd
    from noVis_sim import DroneEnv                     # the reference's own env
    import models
    import train
    np.random.seed(100 + wid); torch.manual_seed(100 + wid)
    env = DroneEnv(row_count=grid, col_count=grid, step_size=1.0, n_anamolous=5, n_drones=n_drones)
    nets = [models.Policy(env.state_size, n_drones, env.action_size, policy_type="MLP") for _ in range(n_drones)]
    opts = [torch.optim.AdamW(nets[i].parameters(), lr=2e-4) for i in range(n_drones)]

    def run(iters):
        args = types.SimpleNamespace(enable_icm=False, policy="MLP", total_steps=T * iters, rollout_steps=T, gamma=0.99, value_coeff=0.5,
                                     entropy_coeff=0.01, grad_acc=1, grad_norm_limit=10.0, save_freq=10 ** 9 + 7, grid_size=grid)
        logging.getLogger().handlers.clear()
        train.train(args, nets, opts, env, env.state_size, n_drones)
    if warmup:
        run(warmup)
    conn.send("ready"); conn.recv()
    t0 = time.perf_counter()
    run(steps)
    conn.send(time.perf_counter() - t0)


def _dql_worker(conn, wid, grid, n_drones, warmup, steps):
    import numpy as np
    import torch
    torch.set_num_threads(1)
    sys.path.insert(0, os.path.join(REF, "DQL_model"))
    path = os.path.join(REF, "DQL_model", "DQLearning.py")
    src = open(path).read()
    head = src[:src.index("\nenv = DroneEnv(")]          # the script's definitions; its module-level run is restated below
    argv, sys.argv = sys.argv, ["DQLearning.py"]
    ns = {"__name__": "_ref_DQLearning"}
    exec(compile(head, path, "exec"), ns)
    sys.argv = argv
    np.random.seed(200 + wid); torch.manual_seed(200 + wid)
    import random
    random.seed(200 + wid)
    env = ns["DroneEnv"](row_count=grid, col_count=grid, step_size=1.0, n_anamolous=5, n_drones=n_drones)
    agent = ns["Agent"](env.state_size, env.action_size, n_drones)
    state = np.reshape(env.reset(), [1, env.state_size])
    pos = np.array(env.n_drones_pos).reshape((n_drones, 2))

    def one():
        nonlocal state, pos
        action = agent.act(state, pos)                                       # DQLearning.py:169
        nstate, reward, done = env.step(action)
        reward = reward if not done else 1000
        nstate = np.reshape(nstate, [1, env.state_size])
        npos = np.array(env.n_drones_pos).reshape((n_drones, 2))
        agent.memorize(state, pos, action, reward, nstate, npos, done)
        state, pos = nstate, npos
        if done:
            state = np.reshape(env.reset(), [1, env.state_size])
            pos = np.array(env.n_drones_pos).reshape((n_drones, 2))
            return
        if len(agent.memory) > 32:
            agent.replay(32)                                                 # DQLearning.py:187-188
    for _ in range(40):                                # fill the memory past batch_size so that every timed step replays
        one()
    for _ in range(warmup):
        one()
    conn.send("ready"); conn.recv()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    conn.send(time.perf_counter() - t0)


def measure(kind, steps, warmup, procs=0, grid=10, n_drones=8, T=25):
    procs = procs or len(os.sched_getaffinity(0))
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):   # one core per copy: the workers import torch after the fork
        os.environ[k] = "1"
    ctx = mp.get_context("fork")
    conns, ps = [], []
    for w in range(procs):
        a, b = ctx.Pipe()
        args = (b, w, grid, n_drones, T, warmup, steps) if kind == "a2c" else (b, w, grid, n_drones, warmup, steps)
        p = ctx.Process(target=_a2c_worker if kind == "a2c" else _dql_worker, args=args, daemon=True)
        p.start(); conns.append(a); ps.append(p)
    for c in conns:
        assert c.recv() == "ready"
    t0 = time.perf_counter()
    for c in conns:
        c.send("go")
    times = [c.recv() for c in conns]
    wall = time.perf_counter() - t0
    for p in ps:
        p.join(timeout=10)
    env_steps = (T if kind == "a2c" else 1) * steps
    return {"kind": kind, "value": procs * env_steps * n_drones / wall, "procs": procs, "steps": steps, "warmup": warmup, "wall_s": wall,
            "slowest_worker_s": max(times), "env_steps_per_step": T if kind == "a2c" else 1, "grid": grid, "n_drones": n_drones,
            "source": "baseline/_ref (unmodified copies of A2C/train.py + models.py + noVis_sim.py / DQL_model/DQLearning.py + noVis_sim.py)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["a2c", "dql"])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--procs", type=int, default=0)
    ap.add_argument("--rollout-steps", type=int, default=25)
    a = ap.parse_args()
    if not available():
        print(json.dumps({"unavailable": "baseline/_ref is not installed (run __graft_entry__.build() where /root/reference exists)"}))
        return
    print(json.dumps(measure(a.kind, a.steps, a.warmup, a.procs, T=a.rollout_steps)))


if __name__ == "__main__":
    main()
