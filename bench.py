#!/usr/bin/env python
"""bench.py — agent-steps/sec of the B200-native swarm env step (BASELINE.json metric), one JSON line on rank 0.

    python bench.py --gpus 1 --steps 200 --warmup 10              # ours, N=1
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 200 --warmup 10                 # one rank per GPU, envs sharded, no data-path collective
    python bench.py --impl reference --steps 3 --warmup 1         # the reference's CPU algorithm (oracle port) on host cores

A "step" is one pass of the hot path (one fused swarm_step launch) over one batch of N envs.  Default workload is
BASELINE.json configs[2] ("cfg3": 65536 envs x 16 drones, 64x64 grid, 5x5 local windows) — the configuration the
1/2/4/8-GPU metric is quoted on; weak scaling (every rank owns --envs envs).  Inputs are resident in HBM for `value`;
`e2e` goes through swarm_step_host with pinned HOST buffers (H2D actions, D2H obs/reward/done every step).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(REPO, "marl-drones_b200"))
sys.path.insert(0, REPO)

WORKLOADS = {
    # name: (envs per GPU, kwargs)   — SURVEY §8d
    "cfg2": (4096, dict(row_count=32, col_count=32, n_drones=4, step_size=1.0, obs_mode="full")),
    "cfg3": (65536, dict(row_count=64, col_count=64, n_drones=16, step_size=1.0, obs_mode="window", window=5)),
    "cfg4": (16384, dict(row_count=10, col_count=10, n_drones=8, step_size=1.0, obs_mode="full")),
    "cfg5a": (1024, dict(row_count=256, col_count=256, n_drones=64, step_size=1.0, obs_mode="full")),
    "cfg5b": (1024, dict(row_count=256, col_count=256, n_drones=64, step_size=1.0, obs_mode="window", window=5)),
}
A2C_WORKLOAD = "cfg4_a2c"   # BASELINE configs[3]: the A2C/train.py rollout loop on the cfg4 env shape, env + policy end to end
DQL_WORKLOAD = "cfg4_dql"   # the DQL_model/DQLearning.py loop (act + env step + memorise + replay) batched on the cfg4 env shape
METRIC = "agent-steps/sec (device-timed)"
UNIT = "agent-steps/s"


def algorithmic_bytes_per_env_step(kw, s_a=1, s_o=4):
    """canonical, layout-independent bytes one env-step must move (1 B per grid cell)."""
    D, RC = kw["n_drones"], kw["row_count"] * kw["col_count"]
    if kw["obs_mode"] == "full":
        return D * s_a + 16 * D + RC + D + RC * s_o + 21
    WW = kw["window"] ** 2
    return D * s_a + 16 * D + 2 * D + D * WW + D * WW * s_o + 21


def load_peak():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic(workload):
    """DRAM bytes per launch of the step kernel from the committed ncu capture (profiles/traffic.json), else None."""
    try:
        return json.load(open(os.path.join(REPO, "profiles", "traffic.json"))).get(workload)
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "samples": len(sm), "reasons": sorted(reasons)}


def cpu_port_throughput(kw, budget_s=12.0, scan_mode=0, threads=0, steps=None):
    """Times the oracle port of the reference algorithm (literal R*C scan per drone, A2C/noVis_sim.py:75-90) on the host
    cores: a bounded sample of the same workload shape.  Returns (agent_steps_per_s, cores, sample description)."""
    import numpy as np
    from oracle.oracle import OracleBatch, lib
    cores = threads or (lib().oracle_max_threads())
    om = 0 if kw["obs_mode"] == "full" else 1
    okw = dict(row_count=kw["row_count"], col_count=kw["col_count"], n_drones=kw["n_drones"], step_size=kw["step_size"],
               obs_mode=om, window=kw.get("window", 5), auto_reset=True, seed=2026, scan_mode=scan_mode, nthreads=cores)
    n_cpu = 8 * cores
    o = OracleBatch(n_cpu, **okw)
    rs = np.random.RandomState(1234)
    acts = rs.randint(0, 4, (16, n_cpu, kw["n_drones"])).astype(np.uint8)
    o.step(acts[0])
    t0 = time.perf_counter(); o.step(acts[1]); dt = max(time.perf_counter() - t0, 1e-6)
    T = steps if steps else max(2, min(20000, int(budget_s / dt)))
    t0 = time.perf_counter()
    for k in range(T):
        o.step(acts[k % 16])
    wall = time.perf_counter() - t0
    v = n_cpu * kw["n_drones"] * T / wall
    sample = (f"{n_cpu} envs x {T} steps of the workload shape, {'literal R*C scan' if scan_mode == 0 else 'bounded window'} "
              f"C port of A2C/noVis_sim.py step (OpenMP, {cores} threads), {wall:.1f} s")
    return v, cores, sample, wall / T


def run_reference(args, kw, n_envs, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (oracle port, all host threads), rank 0 only."""
    if rank != 0:
        return
    steps, warm = max(args.steps, 1), max(args.warmup, 0)
    import numpy as np
    from oracle.oracle import OracleBatch, lib
    cores = lib().oracle_max_threads()
    om = 0 if kw["obs_mode"] == "full" else 1
    n_cpu = 8 * cores
    o = OracleBatch(n_cpu, row_count=kw["row_count"], col_count=kw["col_count"], n_drones=kw["n_drones"],
                    step_size=kw["step_size"], obs_mode=om, window=kw.get("window", 5), auto_reset=True, seed=2026,
                    scan_mode=0, nthreads=cores)
    acts = np.random.RandomState(1234).randint(0, 4, (16, n_cpu, kw["n_drones"])).astype(np.uint8)
    t0 = time.perf_counter(); o.step(acts[0]); dt = max(time.perf_counter() - t0, 1e-6)
    inner = max(1, min(5000, int(120.0 / (steps + warm) / dt)))      # whole run stays within ~2 minutes
    for w in range(warm):
        for k in range(inner):
            o.step(acts[k % 16])
    t0 = time.perf_counter()
    for s in range(steps):
        for k in range(inner):
            o.step(acts[k % 16])
    wall = time.perf_counter() - t0
    v = n_cpu * kw["n_drones"] * inner * steps / wall
    sample = (f"each step = {inner} env-steps of {n_cpu} envs (bounded sample of {args.workload}), literal R*C-scan C port of "
              f"A2C/noVis_sim.py (OpenMP, {cores} threads)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": wall / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": {"workload": args.workload, **{k: v2 for k, v2 in kw.items()}, "envs_per_step": n_cpu},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL's version banner) print to fd 1; the contract is ONE JSON line on stdout.  Everything written to fd 1
    from here on goes to stderr; emit() writes the JSON line to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def run_a2c(args, rank, world, local):
    """BASELINE configs[3] / SURVEY §8f-1: the batched re-expression of A2C/train.py::train() — per outer iteration T =
    --rollout-steps times (16 stacked MlpPolicy nets forward with autograd -> swarm_sample_actions -> fused env step), then the
    bootstrap pass, swarm_nstep_returns, loss, backward, per-drone clip and AdamW.  A bench "step" is ONE outer iteration;
    value = agent-steps/s = N * D * T * iterations / time, env + policy + update end to end, device-timed."""
    n_envs, kw = WORKLOADS["cfg4"]
    kw = dict(kw)
    if args.envs:
        n_envs = args.envs
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps({"impl": "reference", "metric": "agent-steps/sec (env+policy, A2C train loop)", "config": {"workload": A2C_WORKLOAD},
                              "unavailable": "the reference train() is pure Python/PyTorch under /root/reference and cannot travel to the GPU box; "
                                             "measured in the build container: 47.2 agent-steps/s (8 cores, 10x10, 8 drones, MLP; BASELINE.md)"}))
        return
    quiet_stdout()
    import torch
    import torch.distributed as dist
    from swarm_b200 import BatchedDroneEnv
    from swarm_b200.a2c import A2CTrainer
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl ours) needs a B200; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cuda.matmul.allow_tf32 = bool(args.tf32)
    D, T = kw["n_drones"], args.rollout_steps
    K = args.steps if args.steps != 1024 else 10
    W = max(args.warmup if args.warmup != 10 else 3, 3)
    env = BatchedDroneEnv(n_envs, auto_reset=True, seed=2026, env_id_base=rank * n_envs, **kw)
    torch.manual_seed(1234)                                  # identical initial weights on every rank
    tr = A2CTrainer(env, rollout_steps=T, seed=99 + rank)
    for _ in range(W):
        tr.train_iteration()
    torch.cuda.synchronize(dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = env.launch_count
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(K):
        tr.train_iteration()
    t1.record()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = t0.elapsed_time(t1)
    env_launches = env.launch_count - l0
    # e2e: the same loop, reading the loss back to the host every iteration (the step's result); inputs are produced on the
    # device by the env itself, so there is no per-step H2D
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    for _ in range(K):
        float(tr.train_iteration().item())
    torch.cuda.synchronize(dev)
    e2e_ms = (time.perf_counter() - w0) * 1e3
    # rollout only (no autograd, no update): the acting loop of a trained policy
    with torch.no_grad():
        tr.rollout(grad=False)
        torch.cuda.synchronize(dev)
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record()
        for _ in range(K):
            tr.rollout(grad=False)
        r1.record()
        torch.cuda.synchronize(dev)
    roll_ms = r0.elapsed_time(r1)
    if world > 1:
        t = torch.tensor([ms, e2e_ms, roll_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms, roll_ms = (float(x) for x in t.tolist())
    if rank == 0:
        H = kw["row_count"] * kw["col_count"]
        macs = 3 * H * H + H * H + H * 256 + 256 * 128 + 128 * 128 + 128 * 64 + 64 * 32 + 32 * 32   # one MlpPolicy, dense4 aside
        flops_iter = 2.0 * macs * 2 * D * n_envs * (3 * T + 1)        # forward + backward (2x) over T steps, + the bootstrap forward
        agent_steps = float(n_envs) * D * T * K * world
        try:
            peak = float(json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]); src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
        except Exception:
            peak, src = 1500.0, "fallback (B200_PROFILING.md)"
        ach = flops_iter * K / (ms * 1e-3) / 1e12
        line = {
            "metric": "agent-steps/sec (env+policy, A2C train loop, device-timed)", "value": agent_steps / (ms * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32 policy GEMMs" if args.tf32 else "f32 (policy GEMMs via cuBLAS, as the reference's float32)", "data": "synthetic",
            "config": {"workload": A2C_WORKLOAD, "envs_per_gpu": n_envs, "total_envs": n_envs * world, **kw, "rollout_steps": T,
                       "policy": "MLP (A2C/models.py MlpPolicy, actor+critic per drone = 16 nets, random init)", "optimizer": "AdamW lr 2e-4, per-drone clip 10",
                       "step": "one outer iteration of train(): T x (policy fwd + sample + env step) + bootstrap + returns + loss + backward + update",
                       "l2": "activations kept for backward: ~1 GB per rollout step, far beyond the 126 MB L2",
                       "parallelism": f"data-parallel x{world}: envs sharded, one gradient all-reduce per iteration" if world > 1 else "single GPU"},
            "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": None, "peak_source": src,
                         "kernel": "cuBLAS batched SGEMM behind torch.bmm (library GEMMs: this row is a caller of the hot path, not the hot path)",
                         "algorithmic_flops_per_iteration": flops_iter},
            "e2e": {"value": agent_steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4, "steps": K,
                    "ms_per_step": e2e_ms / K, "api": "A2CTrainer.train_iteration() + loss.item()"},
            "rollout_only": {"value": agent_steps / (roll_ms * 1e-3), "unit": UNIT, "ms_per_step": roll_ms / K, "what": "no autograd, no update: policy fwd + sample + env step"},
            "gpu_launches": int(env_launches + K * (T + 1) + K * 11 * (T + 1)), "gpu_launches_note": "swarm_step + swarm_sample_actions per rollout step, swarm_nstep_returns + reset per iteration, swarm_gru_gates_fwd/_bwd + swarm_relu_l2norm_fwd/_bwd + 7 swarm_bias_act per policy call (T + 1 per iteration; torch/cuBLAS kernels not counted)",
            "clocks": clocks,
            "cpu_baseline": {"value": 47.2, "unit": UNIT, "cores": 8, "kind": "reference", "sample": "unmodified A2C/train.py::train(), 250 env-steps, 10x10, 8 drones, MLP, measured in the build container (cannot travel to the GPU box)"},
        }
        emit(line)
    sys.stdout.flush()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    env.close()
    sys.stdout.flush(); sys.stderr.flush()
    os._exit(0)


def run_dql(args, rank, world, local):
    """The batched re-expression of DQL_model/DQLearning.py: a bench "step" is ONE pass of its inner loop for all
    envs — 8 stacked AgentModel nets act, fused env step (DQL reward), memorise, replay on a minibatch of 32 transitions per env
    (the reference's --batch_size per env step) with one AdamW step.  value = N * D * steps / time, device-timed."""
    n_envs, kw = WORKLOADS["cfg4"]
    kw = dict(kw)
    if args.envs:
        n_envs = args.envs
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps({"impl": "reference", "metric": "agent-steps/sec (env+agent, deep-Q loop)", "config": {"workload": DQL_WORKLOAD},
                              "unavailable": "the reference DQLearning.py is pure Python/PyTorch under /root/reference and cannot travel to the GPU box; "
                                             "measured in the build container: 15.9 agent-steps/s (8 cores, 10x10, 8 drones; profiles/r1_dql_reference_cpu.json)"}))
        return
    quiet_stdout()
    import torch
    import torch.distributed as dist
    from swarm_b200 import BatchedDroneEnv
    from swarm_b200.dql import BatchedDQLAgent
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl ours) needs a B200; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cuda.matmul.allow_tf32 = bool(args.tf32)
    D = kw["n_drones"]
    K = args.steps if args.steps != 1024 else 50
    W = max(args.warmup if args.warmup != 10 else 40, 40)      # the memory must pass batch_size (33 steps) before replay runs
    env = BatchedDroneEnv(n_envs, reward_mode="DQL", auto_reset=True, seed=2026, env_id_base=rank * n_envs, **kw)
    torch.manual_seed(1234)                                  # identical initial weights on every rank
    agent = BatchedDQLAgent(env, memory_steps=48, batch_size=32 * n_envs, seed=99 + rank)
    for _ in range(W):
        agent.train_step()
    torch.cuda.synchronize(dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = env.launch_count
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(K):
        agent.train_step()
    t1.record()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = t0.elapsed_time(t1)
    env_launches = env.launch_count - l0
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    for _ in range(K):                                       # e2e: the loss read back to the host every step
        float(agent.train_step()[2].item())
    torch.cuda.synchronize(dev)
    e2e_ms = (time.perf_counter() - w0) * 1e3
    if world > 1:
        t = torch.tensor([ms, e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms = (float(x) for x in t.tolist())
    if rank == 0:
        S = kw["row_count"] * kw["col_count"]
        macs = S * 64 + 2 * D * 64 + 128 * 64 + 64 * 32 + 32 * 4                                      # one AgentModel
        B = agent.batch_size
        flops_step = 2.0 * macs * D * (n_envs + 4 * B)         # act forward on N rows; replay: Q(s) fwd + 2x bwd, Q(s') fwd on B rows
        agent_steps = float(n_envs) * D * K * world
        try:
            peak = float(json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]); src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
        except Exception:
            peak, src = 1500.0, "fallback (B200_PROFILING.md)"
        ach = flops_step * K / (ms * 1e-3) / 1e12
        line = {
            "metric": "agent-steps/sec (env+agent, deep-Q loop, device-timed)", "value": agent_steps / (ms * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32 GEMMs" if args.tf32 else "f32 (GEMMs via cuBLAS, as the reference's float32)", "data": "synthetic",
            "config": {"workload": DQL_WORKLOAD, "envs_per_gpu": n_envs, "total_envs": n_envs * world, **kw, "reward_mode": "DQL",
                       "agent": "DQL_model/DQLearning.py AgentModel x 8 drones (random init), AdamW lr 1e-4, gamma 0.95, epsilon from 1.0 decaying 0.995 per replay",
                       "replay_batch": B, "memory_steps": agent.K,
                       "step": "act (epsilon-greedy) + env step + memorise + replay (targets, loss, backward, one AdamW step)",
                       "l2": "replay activations [8, 524288, 128] float32 = 2 GB per layer, far beyond the 126 MB L2",
                       "parallelism": f"data-parallel x{world}: envs sharded, one gradient all-reduce per replay" if world > 1 else "single GPU"},
            "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": None, "peak_source": src,
                         "kernel": "cuBLAS SGEMM + element-wise torch kernels (library code: this row is a caller of the hot path, not the hot path)",
                         "algorithmic_flops_per_step": flops_step},
            "e2e": {"value": agent_steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4, "steps": K,
                    "ms_per_step": e2e_ms / K, "api": "BatchedDQLAgent.train_step() + loss.item()"},
            "gpu_launches": int(env_launches + K * 16), "gpu_launches_note": "per step: swarm_step + the masked swarm_reset (step cap, also the state copy); swarm_row_standardize + swarm_embed_tanh_l2norm_fwd + 3 swarm_bias_act per net pass (act, Q(s), Q(s')) and one swarm_embed_tanh_l2norm_bwd; torch/cuBLAS kernels not counted",
            "clocks": clocks,
            "cpu_baseline": {"value": 15.9, "unit": UNIT, "cores": 8, "kind": "reference", "sample": "unmodified DQLearning.py Agent + env loop, 60 env-steps with replay(32), 10x10, 8 drones, measured in the build container by tools/dql_ref_cpu.py (cannot travel to the GPU box)"},
        }
        emit(line)
    sys.stdout.flush()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    env.close()
    sys.stdout.flush(); sys.stderr.flush()
    os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1024)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS) + [A2C_WORKLOAD, DQL_WORKLOAD])
    ap.add_argument("--rollout-steps", type=int, default=25, help="cfg4_a2c: env steps per outer iteration (A2C/main.py default)")
    ap.add_argument("--tf32", action="store_true", help="cfg4_a2c: allow TF32 tensor-core GEMMs for the policy (default: fp32, as the reference)")
    ap.add_argument("--obs-dtype", default="float32", choices=["float32", "uint8"],
                    help="observation element type of the device-timed step (default float32 = the reference's; uint8 is a diagnostic line)")
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (default: the workload's)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--sets", type=int, default=4, help="independent env batches rotated so the working set exceeds L2")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--no-graph", action="store_true", help="launch every step from Python instead of replaying a CUDA graph of 32 steps")
    ap.add_argument("--clock-window-s", type=float, default=0.4, help="seconds of identical untimed steps run under the clock sampler right before the timed region")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--stats-mode", default="peer", choices=["peer", "nccl"],
                    help="N > 1: how the per-step episode statistics cross GPUs — 'peer': pushed by the step kernel's last CTA into "
                         "every rank's mailbox over NVLink peer memory (no collective kernel); 'nccl': swarm_stats_allreduce on a side stream")
    ap.add_argument("--stats-every", type=int, default=1, help="N > 1: all-reduce the episode statistics every K steps (0 = never)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if args.workload == A2C_WORKLOAD:
        return run_a2c(args, rank, world, local)
    if args.workload == DQL_WORKLOAD:
        return run_dql(args, rank, world, local)
    n_envs, kw = WORKLOADS[args.workload]
    kw = dict(kw)
    if args.envs:
        n_envs = args.envs
    if args.scaling == "strong":
        n_envs = n_envs // world
    if args.impl == "reference":
        return run_reference(args, kw, n_envs, rank, world)

    quiet_stdout()
    import torch
    import torch.distributed as dist
    from swarm_b200 import BatchedDroneEnv
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl ours) needs a B200; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    D = kw["n_drones"]
    K, W = args.steps, max(args.warmup, 3)
    obs_dt = torch.float32 if args.obs_dtype == "float32" else torch.uint8
    s_o = 4 if args.obs_dtype == "float32" else 1
    envs = [BatchedDroneEnv(n_envs, auto_reset=True, seed=2026, env_id_base=(rank * args.sets + s) * n_envs, obs_dtype=obs_dt, **kw)
            for s in range(args.sets)]
    obs = [torch.empty(e.obs_shape, dtype=obs_dt, device=dev) for e in envs]
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    # 32 pre-generated action tensors: 16 i.i.d. uniform draws followed by their opposites ((a+2) mod 4), so a cycle has no
    # net drift and the drones keep wandering instead of piling up against a wall over thousands of steps.
    half = torch.randint(0, 4, (16, n_envs, D), dtype=torch.uint8, device=dev, generator=gen)
    acts = torch.cat([half, (half + 2) % 4]).contiguous()
    n_act = acts.shape[0]
    stream = torch.cuda.current_stream(dev)
    comm = None
    use_nccl = world > 1 and args.stats_mode == "nccl"
    if world > 1 and not use_nccl:
        ok = 1
        try:
            for e in envs:
                e.enable_peer_stats()
        except Exception as exc:            # CUDA IPC unavailable between these processes: every rank switches to the NCCL form
            print(f"[bench] rank {rank}: peer-memory statistics unavailable ({exc!r}); using --stats-mode nccl", file=sys.stderr)
            ok = 0
        flag = torch.tensor([ok], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            use_nccl = True
            for e in envs:
                try:
                    e.disable_peer_stats()
                except Exception:
                    pass
    side = None
    if use_nccl:
        from swarm_b200 import StatsComm
        comm = StatsComm()
        side = torch.cuda.Stream(dev)       # the all-reduce runs beside the step stream
    snaps = [torch.zeros(8, dtype=torch.int64, device=dev) for _ in range(n_act)]

    def one_step(i, ev=None):
        s = i % args.sets
        cur = torch.cuda.current_stream(dev)
        if ev is not None:
            ev[0].record(cur)
        envs[s].step(acts[i % n_act], out=obs[s])
        if ev is not None:
            ev[1].record(cur)
        if use_nccl and i % max(args.stats_every, 1) == 0:   # the path's only collective: 64-byte episode statistics
            snap = snaps[i % n_act]
            snap.copy_(envs[s].stats, non_blocking=True)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                comm.allreduce(snap)        # swarm_stats_allreduce: ncclAllReduce(SUM, int64[8]) on the library's communicator

    for i in range(W):
        one_step(i)
    if side is not None:
        stream.wait_stream(side)
    torch.cuda.synchronize(dev)

    # CUDA graph of n_act consecutive steps (+ their side-stream all-reduces): the host enqueues one graph per 32 steps
    # instead of 32 launches, so a microsecond-scale step is not bound by Python call overhead.
    graph = None
    if not args.no_graph and n_act % args.sets == 0:
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for j in range(n_act):
                    one_step(j)
                if side is not None:
                    torch.cuda.current_stream(dev).wait_stream(side)
            graph.replay()
            torch.cuda.synchronize(dev)
        except Exception as exc:            # capture unsupported (e.g. a collective that refuses capture): eager launches
            print(f"[bench] CUDA graph capture failed ({exc!r}); falling back to eager launches", file=sys.stderr)
            graph = None
            torch.cuda.synchronize(dev)
    if world > 1:                           # all ranks must agree (each step carries a collective)
        flag = torch.tensor([1 if graph is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            graph = None

    def run_steps(first, count, ev=None):
        """`count` consecutive steps starting at step index `first` (a multiple of n_act): graph replays + eager remainder."""
        k = 0
        if graph is not None and ev is None:
            while count - k >= n_act:
                graph.replay()
                k += n_act
        while k < count:
            one_step(first + k, None if ev is None else ev[k])
            k += 1

    # The timed region of a microsecond-scale step lasts milliseconds, shorter than nvidia-smi's sampling period, so the
    # clock sampler runs over a window of identical untimed steps that leads straight into the K timed ones
    # (continuous load, no idle gap); the window is named in the JSON.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter(); run_steps(0, n_act); torch.cuda.synchronize(dev)
    est = max((time.perf_counter() - t0) / n_act, 5e-6)
    if world > 1:                           # every rank must run the same number of steps
        te = torch.tensor([est], dtype=torch.float64, device=dev)
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        est = float(te.item())
    hold = int(min(40000, max(0, args.clock_window_s / est))) // n_act * n_act
    run_steps(0, hold)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_start.record(stream)
    run_steps(0, K)
    if side is not None:
        stream.wait_stream(side)
    t_end.record(stream)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total = t_start.elapsed_time(t_end)
    launches = K                            # one step kernel per step (graph replays launch the captured kernels)
    # per-launch durations: a separate eager pass with an event pair around every launch (not part of `value`)
    P = min(K, 64)
    pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(P)]
    run_steps(0, P, pairs)
    if side is not None:
        stream.wait_stream(side)
    torch.cuda.synchronize(dev)
    kern_ms = sorted(a.elapsed_time(b) for a, b in pairs)
    pair_avg_ms = sum(kern_ms) / len(kern_ms)
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    # the stream holds nothing but step kernels back to back, so the mean launch-to-launch time over the timed region
    # bounds the kernel's duration from above (it includes the inter-kernel gaps): the conservative roofline numerator
    kern_avg_ms = ms_total / K
    total_agent_steps = float(n_envs) * D * K * world
    value = total_agent_steps / (ms_total * 1e-3)

    # ---- e2e: host buffers through swarm_step_host (H2D actions + D2H obs/reward/done inside the timed region)
    e0 = envs[0]
    a_h = torch.empty((n_act, n_envs, D), dtype=torch.uint8).pin_memory(); a_h.copy_(acts.cpu())
    obs_h = torch.empty(e0.obs_shape, dtype=obs_dt).pin_memory()
    rew_h = torch.empty(n_envs, dtype=torch.float32).pin_memory()
    done_h = torch.empty(n_envs, dtype=torch.uint8).pin_memory()
    stats_h = torch.zeros(8, dtype=torch.int64).pin_memory()
    E = max(1, args.e2e_steps)
    for i in range(3):
        envs[i % args.sets].step_host(a_h[i % n_act], obs_h, rew_h, done_h, stats_h)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    for i in range(E):
        envs[i % args.sets].step_host(a_h[i % n_act], obs_h, rew_h, done_h, stats_h)
    e2e_ms = (time.perf_counter() - w0) * 1e3
    if world > 1:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e_value = float(n_envs) * D * E * world / (e2e_ms * 1e-3)
    h2d = n_envs * D
    d2h = int(obs_h.numel() * s_o + n_envs * 4 + n_envs + 64)
    # the same host-buffer call with uint8 observations (values 0 / 1 / U, pad 255): the reference's obs are small integers
    # (int64 on the host, A2C/noVis_sim.py:168-170) that callers cast to float32 themselves, so a byte per cell loses nothing
    # and quarters the PCIe traffic that bounds e2e.  Reported beside, not instead of, the float32 number.
    e8 = BatchedDroneEnv(n_envs, auto_reset=True, seed=2026, env_id_base=rank * args.sets * n_envs, obs_dtype=torch.uint8, **kw)
    obs8_h = torch.empty(e8.obs_shape, dtype=torch.uint8).pin_memory()
    for i in range(3):
        e8.step_host(a_h[i % n_act], obs8_h, rew_h, done_h, stats_h)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    for i in range(E):
        e8.step_host(a_h[i % n_act], obs8_h, rew_h, done_h, stats_h)
    e8_ms = (time.perf_counter() - w0) * 1e3
    if world > 1:
        t = torch.tensor([e8_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e8_ms = float(t.item())
    e8_value = float(n_envs) * D * E * world / (e8_ms * 1e-3)
    e8.close()

    stats_global = None
    if world > 1 and not use_nccl:          # what the mailboxes hold after the run: totals of ALL ranks for env batch 0
        torch.cuda.synchronize(dev)
        dist.barrier()
        local = envs[0].stats.clone()
        tot = local.clone()
        dist.all_reduce(tot)                # checker only (outside every timed region): NCCL sum of the local totals
        envs[0].push_stats()                # the in-step feed lags by one launch: publish the final totals explicitly
        torch.cuda.synchronize(dev)
        dist.barrier()
        stats_global = envs[0].global_stats(min_pushes=1).tolist()
        assert stats_global == tot.tolist(), (stats_global, tot.tolist())
    if rank == 0:
        peak, peak_src = load_peak()
        bytes_per_launch = algorithmic_bytes_per_env_step(kw, s_o=s_o) * n_envs
        achieved = bytes_per_launch / (kern_avg_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "u2 grid codes / f32 positions+obs (integer rewards)", "data": "synthetic",
            "config": {"workload": args.workload, "envs_per_gpu": n_envs, "total_envs": n_envs * world, **kw,
                       "obs_dtype": args.obs_dtype, "auto_reset": True, "reward_mode": "A2C", "reset_rng": "philox4x32-10",
                       "actions": "uint8 uniform{0..3} (16 draws + their opposites, zero net drift), pre-generated on device",
                       "l2": f"{args.sets} independent env batches rotated; working set {args.sets * (e0._obs.numel() * s_o + n_envs * 1100) / 1e6:.0f} MB > 126 MB L2",
                       "parallelism": f"env-sharded x{world}, no data-path collective; per-step episode statistics: " +
                                      ("published by the step kernel into every rank's mailbox over NVLink peer memory (posted stores at kernel start, no collective kernel)"
                                       if not use_nccl else f"swarm_stats_allreduce (NCCL) on a side stream every {max(args.stats_every, 1)} step(s)")},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": load_traffic(args.workload) if args.obs_dtype == "float32" else None, "peak_source": peak_src, "kernel": "swarm::step_warp_kernel (swarm::step_cta_kernel for cfg5a/cfg5b)",
                         "kernel_ms_avg": kern_avg_ms, "kernel_ms_event_pairs_avg": pair_avg_ms, "kernel_ms_event_pairs_median": kern_ms[len(kern_ms) // 2],
                         "launch": "CUDA graph of 32 steps" if graph is not None else "eager",
                         "algorithmic_bytes_per_launch": bytes_per_launch,
                         "algorithmic_bytes_per_agent_step": algorithmic_bytes_per_env_step(kw, s_o=s_o) / D},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": E, "ms_per_step": e2e_ms / E, "api": "swarm_step_host (pinned host buffers, full obs D2H)"},
            "e2e_uint8_obs": {"value": e8_value, "unit": UNIT, "h2d_bytes_per_step": h2d,
                              "d2h_bytes_per_step": int(obs8_h.numel() + n_envs * 4 + n_envs + 64), "steps": E, "ms_per_step": e8_ms / E},
            "stats_global_env_batch0": stats_global,
            "gpu_launches": launches, "clocks": dict(clocks, window=f"{hold} untimed steps + the {K} timed steps, continuous load"),
        }
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample, _ = cpu_port_throughput(kw, budget_s=12.0, scan_mode=0)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
            v2, _, sample2, _ = cpu_port_throughput(kw, budget_s=4.0, scan_mode=1)
            line["cpu_baseline_windowed"] = {"value": v2, "unit": UNIT, "cores": cores, "kind": "port (ours: bounded-window scan, not the reference's algorithm)", "sample": sample2}
        emit(line)
    sys.stdout.flush()
    # teardown: release the captured graph (it holds NCCL work) before the communicator, and leave without running the
    # interpreter's module teardown, which can dead-lock on a communicator that was used under graph capture
    graph = None
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(dev)
    for e in envs:
        e.close()
    if comm is not None:
        comm.close()
    sys.stdout.flush(); sys.stderr.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
