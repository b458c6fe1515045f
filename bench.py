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


def real_bytes_per_env_step(kw, s_a=1, s_o=4):
    """DRAM bytes one env-step moves with the resident layout of DESIGN.md §3 (1 bit per cell, 32-byte anomaly list, float2
    positions read + written, int4 meta read + written) — the layout-dependent counterpart of algorithmic_bytes_per_env_step."""
    D, RC = kw["n_drones"], kw["row_count"] * kw["col_count"]
    words = ((RC + 31) // 32 + 3) // 4 * 4
    state = words * 4 + 32 + 16 * D + 32 + D * s_a + 5
    if kw["obs_mode"] == "full":
        return state + RC * s_o
    return state + D * kw["window"] ** 2 * s_o


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


def host_cores():
    """Cores this process may run on.  torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arms ignore it and use every
    core of the box (the N >= 2 reference arms ran on one thread)."""
    if os.environ.get("SWARM_BENCH_CORES"):      # test knob (tests/test_bench_contract.py): a small worker pool for the slow arms
        return max(1, int(os.environ["SWARM_BENCH_CORES"]))
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_port_throughput(kw, budget_s=12.0, scan_mode=0, threads=0, steps=None):
    """Times the oracle port of the reference algorithm (literal R*C scan per drone, A2C/noVis_sim.py:75-90) on the host
    cores: a bounded sample of the same workload shape.  Returns (agent_steps_per_s, cores, sample description)."""
    import numpy as np
    from oracle.oracle import OracleBatch
    cores = threads or host_cores()
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


def python_reference_throughput(kw, steps, warmup, budget_s):
    """The UNMODIFIED reference DroneEnv (baseline/_ref/A2C/noVis_sim.py) on every host core: P worker processes, no per-step
    IPC (baseline/ref_runner.py; SURVEY §8(d) primary CPU baseline).  Runs in a child interpreter so that the workers are
    forked from a process that never touched CUDA.  Returns the runner's dict, or None when baseline/_ref is not installed."""
    runner = os.path.join(REPO, "baseline", "ref_runner.py")
    cmd = [sys.executable, runner, "--steps", str(steps), "--warmup", str(warmup), "--budget-s", str(budget_s),
           "--procs", str(host_cores()), "--row_count", str(kw["row_count"]), "--col_count", str(kw["col_count"]),
           "--n_drones", str(kw["n_drones"]), "--step_size", str(kw["step_size"])]
    env = dict(os.environ)
    env.pop("OMP_NUM_THREADS", None)
    try:
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=budget_s * 4 + 120, env=env)
        d = json.loads(p.stdout.strip().splitlines()[-1])
    except Exception as exc:
        print(f"[bench] python reference runner failed: {exc!r}", file=sys.stderr)
        return None
    return None if "unavailable" in d else d


def reference_sample_text(r, workload):
    return (f"each step = {r['inner']} env-steps of {r['procs'] * r['envs_per_proc']} envs (bounded sample of {workload}: the shape is the "
            f"workload's, the env count is not — CPU throughput does not depend on it), UNMODIFIED A2C/noVis_sim.py DroneEnv from "
            f"baseline/_ref, {r['procs']} worker processes, reset on done, random actions, {r['wall_s']:.1f} s")


def run_reference(args, kw, n_envs, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on every host core, rank 0 only.  Primary: the
    unmodified Python DroneEnv installed under baseline/_ref (kind "reference"); the oracle's C/OpenMP port of the same
    algorithm (kind "port", far faster than the reference itself) is reported beside it as `cpu_port`, and is the fallback
    when baseline/_ref is absent."""
    if rank != 0:
        return
    steps, warm = max(args.steps, 1), max(args.warmup, 0)
    cores = host_cores()
    cfg = {"workload": args.workload, **{k: v2 for k, v2 in kw.items()}}
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic"}
    r = python_reference_throughput(kw, steps, warm, budget_s=min(60.0, 10.0 + 2.5 * (steps + warm)))
    if r is not None:
        port_v, _, port_sample, _ = cpu_port_throughput(kw, budget_s=8.0, scan_mode=0, threads=cores)
        v = r["value"]
        cfg.update(total_envs=r["procs"] * r["envs_per_proc"], env_steps_per_step=r["inner"],
                   note="bounded sample: P = host cores envs of the workload's shape, not the workload's env count")
        line.update(value=v, ms_per_step=r["wall_s"] / steps * 1e3, config=cfg,
                    cpu_baseline={"value": v, "unit": UNIT, "cores": r["procs"], "kind": "reference",
                                  "sample": reference_sample_text(r, args.workload)},
                    cpu_port={"value": port_v, "unit": UNIT, "cores": cores, "kind": "port", "sample": port_sample})
    else:
        import numpy as np
        from oracle.oracle import OracleBatch
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
        cfg.update(total_envs=n_cpu, env_steps_per_step=inner,
                   note="baseline/_ref not installed: the oracle's C port of the reference algorithm stands in")
        sample = (f"each step = {inner} env-steps of {n_cpu} envs (bounded sample of {args.workload}), literal R*C-scan C port of "
                  f"A2C/noVis_sim.py (OpenMP, {cores} threads)")
        line.update(value=v, ms_per_step=wall / steps * 1e3, config=cfg,
                    cpu_baseline={"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample})
    line["e2e"] = {"value": line["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    print(json.dumps(line))


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


def run_trainer_reference(args, rank, kind, workload, metric, kw):
    """--impl reference for the caller-row workloads: the UNMODIFIED reference loop (A2C/train.py::train() with models.Policy,
    or the DQL_model/DQLearning.py loop) from baseline/_ref, one copy per host core (baseline/ref_trainers.py), rank 0 only.
    A step is one outer iteration of train() on ONE env (the reference is batch-1) / one pass of the deep-Q loop; the step
    count is capped so that the run stays within a few minutes."""
    if rank != 0:
        return
    steps = max(1, min(args.steps if args.steps != 1024 else 10, 12 if kind == "a2c" else 60))
    warm = max(0, min(args.warmup, 2))
    cmd = [sys.executable, os.path.join(REPO, "baseline", "ref_trainers.py"), kind, "--steps", str(steps), "--warmup", str(warm),
           "--procs", str(host_cores()), "--rollout-steps", str(args.rollout_steps)]
    env = dict(os.environ)
    env.pop("OMP_NUM_THREADS", None)
    try:
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=1500, env=env)
        r = json.loads(p.stdout.strip().splitlines()[-1])
    except Exception as exc:
        r = {"unavailable": f"reference trainer runner failed: {exc!r}"}
    if "unavailable" in r:
        print(json.dumps({"impl": "reference", "metric": metric, "config": {"workload": workload}, "unavailable": r["unavailable"]}))
        return
    sample = (f"{r['procs']} independent copies (one per host core) of the UNMODIFIED reference loop from baseline/_ref, batch 1 each, "
              f"{r['grid']}x{r['grid']}, {r['n_drones']} drones, {steps} steps of {r['env_steps_per_step']} env-step(s), {r['wall_s']:.1f} s")
    v = r["value"]
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": r["wall_s"] / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": workload, **kw, "total_envs": r["procs"], "rollout_steps": args.rollout_steps,
                                         "note": "bounded sample: the reference loop is batch-1, so P = host cores envs, not the workload's 16 384"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": r["procs"], "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def teardown(dev, world, envs, comm=None, graphs=()):
    """Orderly exit (no os._exit: the interpreter's exit handlers must run so that the driver's hook sees which .so files
    this process mapped).  Order matters: captured graphs first (they reference the env buffers and, in --stats-mode nccl,
    NCCL work), then a barrier so that no peer is still pushing statistics into this rank's mailbox, then the peer mappings,
    the library communicator, the env handles, and the process group last."""
    import torch
    import torch.distributed as dist
    sys.stdout.flush()
    for g in graphs:
        if g is not None:
            g.reset()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(dev)
    for e in envs:
        try:
            e.disable_peer_stats()
        except Exception:
            pass
    if world > 1:
        dist.barrier()
    if comm is not None:
        comm.close()
    for e in envs:
        e.close()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush(); sys.stderr.flush()


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
        return run_trainer_reference(args, rank, "a2c", A2C_WORKLOAD, "agent-steps/sec (env+policy, A2C train loop, device-timed)", kw)
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
    teardown(dev, world, [env])


def run_dql(args, rank, world, local):
    """The batched re-expression of DQL_model/DQLearning.py: a bench "step" is ONE pass of its inner loop for all
    envs — 8 stacked AgentModel nets act, fused env step (DQL reward), memorise, replay on a minibatch of 32 transitions per env
    (the reference's --batch_size per env step) with one AdamW step.  value = N * D * steps / time, device-timed."""
    n_envs, kw = WORKLOADS["cfg4"]
    kw = dict(kw)
    if args.envs:
        n_envs = args.envs
    if args.impl == "reference":
        return run_trainer_reference(args, rank, "dql", DQL_WORKLOAD, "agent-steps/sec (env+agent, deep-Q loop, device-timed)", kw)
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
    teardown(dev, world, [env])


class StepLoop:
    """`sets` independent env batches of n_envs envs each (rotated so that the working set exceeds L2), 32 pre-generated action
    tensors, and the timed loop: K consecutive steps as CUDA-graph replays — floor(K / 32) replays of a 32-step graph plus ONE
    replay of a (K mod 32)-step graph — so the timed region holds nothing but step kernels whatever --steps is."""

    N_ACT = 32

    def __init__(self, torch, dev, n_envs, kw, sets, obs_dt, id_base, rank, use_graph=True, id_stride=None):
        from swarm_b200 import BatchedDroneEnv
        self.torch, self.dev, self.n_envs, self.sets, self.kw = torch, dev, n_envs, sets, kw
        D = kw["n_drones"]
        self.envs = [BatchedDroneEnv(n_envs, auto_reset=True, seed=2026, env_id_base=id_base + s * (id_stride or n_envs), obs_dtype=obs_dt, **kw)
                     for s in range(sets)]
        self.obs = [torch.empty(e.obs_shape, dtype=obs_dt, device=dev) for e in self.envs]
        gen = torch.Generator(device=dev).manual_seed(1234 + rank)
        # 32 pre-generated action tensors: 16 i.i.d. uniform draws followed by their opposites ((a+2) mod 4), so a cycle has no
        # net drift and the drones keep wandering instead of piling up against a wall over thousands of steps.
        half = torch.randint(0, 4, (16, n_envs, D), dtype=torch.uint8, device=dev, generator=gen)
        self.acts = torch.cat([half, (half + 2) % 4]).contiguous()
        self.use_graph = use_graph and self.N_ACT % sets == 0
        self.graphs = {}            # n steps -> CUDAGraph of steps 0 .. n-1
        self.hook = None            # optional per-step callable(i, env) enqueued after the step (NCCL statistics form)

    def one_step(self, i, ev=None):
        s = i % self.sets
        if ev is not None:
            ev[0].record(self.torch.cuda.current_stream(self.dev))
        self.envs[s].step(self.acts[i % self.N_ACT], out=self.obs[s])
        if ev is not None:
            ev[1].record(self.torch.cuda.current_stream(self.dev))
        if self.hook is not None:
            self.hook(i, self.envs[s])

    def graph_for(self, n, tail=None):
        """CUDA graph of steps 0 .. n-1 (captured once, replayed once as warm-up).  None when capture is off / impossible."""
        torch = self.torch
        if not self.use_graph or n <= 0:
            return None
        if n not in self.graphs:
            try:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    for j in range(n):
                        self.one_step(j)
                    if tail is not None:
                        tail()
                g.replay()
                torch.cuda.synchronize(self.dev)
                self.graphs[n] = g
            except Exception as exc:    # capture unsupported (e.g. a collective that refuses capture): eager launches
                print(f"[bench] CUDA graph capture of {n} steps failed ({exc!r}); falling back to eager launches", file=sys.stderr)
                self.use_graph = False
                torch.cuda.synchronize(self.dev)
                return None
        return self.graphs[n]

    def timed_graph(self, n, tail=None):
        """A graph of [record e0][steps 0 .. n-1][record e1] with EXTERNAL timing events (cudaEventRecordExternal nodes): the
        elapsed time between the two nodes is the device time of exactly n steps, without the host's graph-launch latency that
        an event pair around a single short replay would include.  Returns (graph, e0, e1) or None."""
        torch = self.torch
        if not self.use_graph or n <= 0:
            return None
        try:
            e0 = torch.cuda.Event(enable_timing=True, external=True)
            e1 = torch.cuda.Event(enable_timing=True, external=True)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                e0.record(torch.cuda.current_stream(self.dev))
                for j in range(n):
                    self.one_step(j)
                if tail is not None:
                    tail()
                e1.record(torch.cuda.current_stream(self.dev))
            g.replay()
            torch.cuda.synchronize(self.dev)
            if not (e0.elapsed_time(e1) > 0):
                raise RuntimeError("external events did not time")
            self.graphs[("timed", n)] = g
            return g, e0, e1
        except Exception as exc:
            print(f"[bench] timed graph with external events unavailable ({exc!r}); timing around the replay", file=sys.stderr)
            torch.cuda.synchronize(self.dev)
            return None

    def plan(self, count, tail=None):
        """(full replays, remainder) of `count` steps and the graphs they need; eager when graphs are off."""
        full, rem = divmod(count, self.N_ACT)
        gf = self.graph_for(self.N_ACT, tail) if full else None
        gr = self.graph_for(rem, tail) if rem else None
        if (full and gf is None) or (rem and gr is None):
            return None
        return full, gf, gr

    def run(self, count, plan=None, ev=None):
        """`count` consecutive steps starting at step index 0 (mod 32)."""
        if plan is not None and ev is None:
            full, gf, gr = plan
            for _ in range(full):
                gf.replay()
            if gr is not None:
                gr.replay()
            return
        for k in range(count):
            self.one_step(k, None if ev is None else ev[k])

    def describe(self, count, plan):
        if plan is None:
            return "eager (one swarm_step call per step from Python)"
        full, rem = divmod(count, self.N_ACT)
        parts = ([f"{full} x {self.N_ACT}-step graph"] if full else []) + ([f"1 x {rem}-step graph"] if rem else [])
        return "CUDA graph replay: " + " + ".join(parts)

    def close(self):
        for g in self.graphs.values():
            g.reset()
        self.graphs = {}


def time_steps(torch, dist, dev, world, loop, K, plan, tail_sync=None, timed=None):
    """Device time of exactly K steps, bracketed by barrier + synchronize, max over ranks (ms).  `timed` = (graph, e0, e1) of
    StepLoop.timed_graph: the events are nodes of the replayed graph itself (used when the whole region is ONE short graph)."""
    stream = torch.cuda.current_stream(dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    if timed is not None:
        timed[0].replay()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        ms = timed[1].elapsed_time(timed[2])
    else:
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(stream)
        loop.run(K, plan)
        if tail_sync is not None:
            tail_sync()
        t1.record(stream)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        ms = t0.elapsed_time(t1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def measure_e2e(torch, dist, dev, world, env, acts_host, obs_dt, E):
    """E steps through swarm_step_host: pinned HOST buffers in and out, H2D actions + D2H obs / reward / done / statistics and a
    stream synchronize inside every call.  Returns (ms total as max over ranks, d2h bytes per step)."""
    n_envs = env.num_envs
    obs_h = torch.empty(env.obs_shape, dtype=obs_dt).pin_memory()
    rew_h = torch.empty(n_envs, dtype=torch.float32).pin_memory()
    done_h = torch.empty(n_envs, dtype=torch.uint8).pin_memory()
    stats_h = torch.zeros(8, dtype=torch.int64).pin_memory()
    n_act = acts_host.shape[0]
    for i in range(3):
        env.step_host(acts_host[i % n_act], obs_h, rew_h, done_h, stats_h)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    for i in range(E):
        env.step_host(acts_host[i % n_act], obs_h, rew_h, done_h, stats_h)
    ms = (time.perf_counter() - w0) * 1e3
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    d2h = int(obs_h.numel() * obs_h.element_size() + n_envs * 4 + n_envs + 64)
    return ms, d2h


def measure_rollout(torch, dist, dev, world, envs, acts, T, D, per, bytes_env, peak, steps_target, use_graph):
    """swarm_rollout: T env steps per launch from actions[T, N, D], observations of every step written to obs[T, N, ...]
    (coverage, positions and meta stay on chip between the T steps).  One CUDA graph of 2 rollouts per env batch, replayed;
    device-timed, max over ranks."""
    acts_T = acts[:T].contiguous() if T <= acts.shape[0] else acts.repeat((T + acts.shape[0] - 1) // acts.shape[0], 1, 1)[:T].contiguous()
    obs_T = torch.empty((T,) + tuple(envs[0].obs_shape), dtype=envs[0].obs_dtype, device=dev)
    rew_T = torch.empty((T, per), dtype=torch.float32, device=dev)
    done_T = torch.empty((T, per), dtype=torch.uint8, device=dev)

    def rollouts():
        for _ in range(2):
            for e in envs:
                e.rollout(acts_T, obs_out=obs_T, reward_out=rew_T, done_out=done_T)
    rollouts()
    torch.cuda.synchronize(dev)
    g = None
    if use_graph:
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                rollouts()
            g.replay()
            torch.cuda.synchronize(dev)
        except Exception as exc:
            print(f"[bench] rollout graph capture failed ({exc!r}); eager", file=sys.stderr)
            g = None
    if world > 1:
        flag = torch.tensor([1 if g is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            g = None
    steps_per = 2 * len(envs) * T
    reps = max(1, steps_target // steps_per)
    stream = torch.cuda.current_stream(dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    for _ in range(reps):
        g.replay() if g is not None else rollouts()
    t1.record(stream)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    msr = t0.elapsed_time(t1)
    if world > 1:
        t = torch.tensor([msr], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        msr = float(t.item())
    usr = msr / (reps * steps_per) * 1e3
    out = {"us_per_step": usr, "value": float(per) * world * D / (usr * 1e-6), "unit": UNIT,
           "roofline_frac": bytes_env * per / (usr * 1e-6) / 1e9 / peak, "steps": reps * steps_per, "rollout_T": T,
           "launch": ("CUDA graph of " if g is not None else "eager: ") + f"{2 * len(envs)} swarm_rollout(T={T}) launches, obs of every step written",
           "obs_buffer_mb": obs_T.numel() * obs_T.element_size() / 1e6}
    if g is not None:
        g.reset()
    del obs_T
    return out


def copy_same_bytes_us(torch, dev, nbytes, n=32, reps=8):
    """A plain device copy that moves as many bytes (read + write) as one step of the workload really moves, under the same
    conditions as `value`: CUDA graph of n copies over 4 rotating buffer pairs (> L2), back to back.  What the memory system
    gives a launch of THIS size — the measured copy peak of MEASURED_PEAKS.json is for 4 GB copies, whose ramp and tail are
    amortised away."""
    half = max(1 << 20, int(nbytes) // 2 // 16 * 16)
    src = [torch.empty(half, dtype=torch.uint8, device=dev) for _ in range(4)]
    dst = [torch.empty(half, dtype=torch.uint8, device=dev) for _ in range(4)]
    for i in range(4):
        dst[i].copy_(src[i])
    torch.cuda.synchronize(dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(n):
            dst[i % 4].copy_(src[i % 4])
    g.replay()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(reps):
        g.replay()
    t1.record()
    torch.cuda.synchronize(dev)
    us = t0.elapsed_time(t1) / (n * reps) * 1e3
    g.reset()
    return us, 2 * half


def launch_floor_us(torch, dev, n=32, reps=20):
    """Step-to-step time of a CUDA graph of n dependent launches of a one-element kernel: what ANY one-launch-per-step loop
    pays per step on this box before doing work (context for the small workloads' roofline fractions)."""
    x = torch.zeros(1, device=dev)
    for _ in range(3):
        x.add_(1.0)
    torch.cuda.synchronize(dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            x.add_(1.0)
    g.replay()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(reps):
        g.replay()
    t1.record()
    torch.cuda.synchronize(dev)
    us = t0.elapsed_time(t1) / (n * reps) * 1e3
    g.reset()
    return us


def measure_strong(torch, dist, dev, world, rank, kw, total_envs, args, value_1gpu, ms_step_1gpu, weak_loop, rollout_1gpu):
    """Strong scaling of the workload: `total_envs` envs sharded over the ranks (contiguous env-id ranges, no data-path
    collective).  Two launch forms, both timed on the device as the max over ranks:
      single-step: one swarm_step launch per env step, 32 of them per CUDA graph (the policy-in-the-loop form);
      rollout:     swarm_rollout — T env steps per launch from actions[T, N, D], observations of every step written to
                   obs[T, N, D, W*W] (grid / positions / meta stay on chip between the T steps).
    At 1 GPU the single-step form IS the weak-scaling measurement (same envs per GPU), so its numbers are reused."""
    D = kw["n_drones"]
    per = total_envs // world
    bytes_env = algorithmic_bytes_per_env_step(kw)
    peak, _ = load_peak()
    out = {"total_envs": per * world, "envs_per_gpu": per, "scaling": "strong"}
    Ks = max(32, args.strong_steps // 32 * 32)
    if world == 1:
        loop, own = weak_loop, False
        sets_s = weak_loop.sets
        ms = ms_step_1gpu * Ks
    else:
        sets_s = 2
        while sets_s < 16 and sets_s * per * 2437 < 300e6:
            sets_s *= 2
        # env ids: batch s of rank g = [s * total + g * per, s * total + (g + 1) * per) — contiguous ranges per rank
        loop, own = StepLoop(torch, dev, per, kw, sets_s, torch.float32, rank * per, rank, use_graph=not args.no_graph,
                             id_stride=per * world), True
        for i in range(4):
            loop.one_step(i)
        torch.cuda.synchronize(dev)
        plan = loop.plan(Ks)
        if dist.is_initialized():
            flag = torch.tensor([1 if plan is not None else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 0:
                plan = None
        loop.run(Ks, plan)                  # warm
        ms = time_steps(torch, dist, dev, world, loop, Ks, plan)
        out["single_step_launch"] = loop.describe(Ks, plan)
    us = ms / Ks * 1e3
    out["single_step"] = {"us_per_step": us, "value": float(per) * world * D / (us * 1e-6), "unit": UNIT,
                          "roofline_frac": bytes_env * per / (us * 1e-6) / 1e9 / peak, "steps": Ks, "env_batches_rotated": sets_s}
    out["rollout"] = rollout_1gpu if world == 1 else measure_rollout(torch, dist, dev, world, loop.envs, loop.acts, max(1, args.rollout_T), D,
                                                                     per, bytes_env, peak, Ks, not args.no_graph)
    if own:
        loop.close()
        for e in loop.envs:
            e.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1024)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS) + [A2C_WORKLOAD, DQL_WORKLOAD])
    ap.add_argument("--rollout-steps", type=int, default=25, help="cfg4_a2c: env steps per outer iteration (A2C/main.py default)")
    ap.add_argument("--tf32", action="store_true", help="cfg4_a2c: allow TF32 tensor-core GEMMs for the policy (default: fp32, as the reference)")
    ap.add_argument("--obs-dtype", default="float32", choices=["float32", "uint8", "packed2"],
                    help="observation element type of the device-timed step (default float32 = the reference's; uint8 and packed2 — "
                         "5x5 windows as one 64-bit word per drone — are diagnostic lines)")
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (default: the workload's)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--sets", type=int, default=4, help="independent env batches rotated so the working set exceeds L2")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--no-graph", action="store_true", help="launch every step from Python instead of replaying CUDA graphs")
    ap.add_argument("--clock-window-s", type=float, default=0.4, help="seconds of identical untimed steps run under the clock sampler right before the timed region")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the extra strong-scaling measurement of the default cfg3 line")
    ap.add_argument("--strong-steps", type=int, default=256)
    ap.add_argument("--no-rollout", action="store_true", help="skip the swarm_rollout(T) measurement of the workload")
    ap.add_argument("--rollout-T", type=int, default=16, help="strong scaling: env steps per swarm_rollout launch")
    ap.add_argument("--stats-mode", default="peer", choices=["peer", "nccl"],
                    help="N > 1: how the per-step episode statistics cross GPUs — 'peer': pushed by the step kernel into "
                         "every rank's mailbox over NVLink peer memory (no collective kernel); 'nccl': swarm_stats_allreduce on a side stream")
    ap.add_argument("--stats-every", type=int, default=1, help="N > 1: all-reduce the episode statistics every K steps (0 = never)")
    ap.add_argument("--dwell", action="store_true",
                    help="dwell goals of SURVEY App. B.2 (swarmsim-2d/run_env.py:46-47): a cell is mapped only after 10 (normal) / 35 "
                         "(anomalous) drone-ticks on it; the second run SURVEY §8(d) asks of cfg5.  The CPU baselines beside it stay "
                         "the noVis env, which has no dwell")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if args.workload == A2C_WORKLOAD:
        return run_a2c(args, rank, world, local)
    if args.workload == DQL_WORKLOAD:
        return run_dql(args, rank, world, local)
    n_envs, kw = WORKLOADS[args.workload]
    kw = dict(kw)
    if args.dwell:
        kw.update(dwell_normal=10, dwell_anom=35)
    workload_envs = n_envs
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
    K, W = max(args.steps, 1), max(args.warmup, 3)
    obs_dt = {"float32": torch.float32, "uint8": torch.uint8, "packed2": torch.int64}[args.obs_dtype]
    s_o = {"float32": 4, "uint8": 1, "packed2": 8.0 / 25.0}[args.obs_dtype]      # bytes per obs cell
    loop = StepLoop(torch, dev, n_envs, kw, args.sets, obs_dt, rank * args.sets * n_envs, rank, use_graph=not args.no_graph)
    envs, n_act = loop.envs, loop.N_ACT
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
    side, tail = None, None
    if use_nccl:
        from swarm_b200 import StatsComm
        comm = StatsComm()
        side = torch.cuda.Stream(dev)       # the all-reduce runs beside the step stream
        snaps = [torch.zeros(8, dtype=torch.int64, device=dev) for _ in range(n_act)]

        def hook(i, env):                   # the path's only collective: 64-byte episode statistics
            if i % max(args.stats_every, 1) == 0:
                cur = torch.cuda.current_stream(dev)
                snap = snaps[i % n_act]
                snap.copy_(env.stats, non_blocking=True)
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    comm.allreduce(snap)    # swarm_stats_allreduce: ncclAllReduce(SUM, int64[8]) on the library's communicator
        loop.hook = hook

        def tail():
            torch.cuda.current_stream(dev).wait_stream(side)

    for i in range(W):                      # W eager warm-up steps (the graphs get one warm-up replay each on top)
        loop.one_step(i)
    if tail is not None:
        tail()
    torch.cuda.synchronize(dev)

    plan = loop.plan(K, tail)
    hold_plan = loop.plan(n_act, tail)
    if world > 1:                           # all ranks must agree (in NCCL mode each step carries a collective)
        flag = torch.tensor([1 if (plan is not None and hold_plan is not None) else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            plan = hold_plan = None

    # K <= 32: the whole timed region is one short graph — time it between event nodes INSIDE the graph, so the host's
    # graph-launch latency (~15 us against a 0.6 ms region at the driver's --steps 20) is not booked as kernel time
    timed = loop.timed_graph(K, tail) if (plan is not None and K <= n_act) else None
    if world > 1:
        flag = torch.tensor([1 if timed is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            timed = None
    # The timed region of a microsecond-scale step lasts milliseconds, shorter than nvidia-smi's sampling period, so the
    # clock sampler runs over a window of identical untimed steps that leads straight into the K timed ones
    # (continuous load, no idle gap); the window is named in the JSON.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter(); loop.run(n_act, hold_plan); torch.cuda.synchronize(dev)
    est = max((time.perf_counter() - t0) / n_act, 5e-6)
    if world > 1:                           # every rank must run the same number of steps
        te = torch.tensor([est], dtype=torch.float64, device=dev)
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        est = float(te.item())
    hold = int(min(40000, max(0, args.clock_window_s / est))) // n_act
    for _ in range(hold):
        loop.run(n_act, hold_plan)
    ms_total = time_steps(torch, dist, dev, world, loop, K, plan, tail, timed)
    clocks = sampler.stop() if rank == 0 else None
    launches = K                            # one step kernel per step (graph replays launch the captured kernels)
    # a second, longer graph-replayed region (diagnostic: shows that `value` does not depend on --steps)
    K_long = 1024
    ms_long = time_steps(torch, dist, dev, world, loop, K_long, loop.plan(K_long, tail) if plan is not None else None, tail) if K != K_long else ms_total
    # per-launch durations: a separate EAGER pass with an event pair around every launch (not part of `value`; includes the
    # launch gaps Python leaves between kernels)
    P = min(K, 64)
    pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(P)]
    loop.run(P, None, pairs)
    if tail is not None:
        tail()
    torch.cuda.synchronize(dev)
    kern_ms = sorted(a.elapsed_time(b) for a, b in pairs)
    pair_avg_ms = sum(kern_ms) / len(kern_ms)
    # under graph replay the stream holds nothing but step kernels back to back, so the mean step-to-step time over the timed
    # region bounds the kernel's duration from above (it includes the inter-kernel gaps): the conservative roofline numerator
    kern_avg_ms = ms_total / K
    total_agent_steps = float(n_envs) * D * K * world
    value = total_agent_steps / (ms_total * 1e-3)

    # ---- the same workload through swarm_rollout (T steps per launch) and the box's per-launch floor: context for the
    # one-launch-per-step value above, most of all for the small workloads (cfg2 / cfg4 / cfg5b move ~10 MB per step)
    rollout = None
    if not args.no_rollout:
        peak_r, _ = load_peak()
        rollout = measure_rollout(torch, dist, dev, world, envs, loop.acts, max(1, args.rollout_T), D, n_envs,
                                  algorithmic_bytes_per_env_step(kw, s_o=s_o), peak_r, 256, not args.no_graph)
    # ---- strong scaling of the same workload (BASELINE configs[2]: "65536 envs ... sharded over 1/2/4/8"): the workload's env
    # count divided over the ranks, contiguous env-id ranges, measured next to the weak-scaling value above
    strong = None
    if args.workload == "cfg3" and args.scaling == "weak" and not args.no_strong and not args.envs and args.obs_dtype == "float32":
        strong = measure_strong(torch, dist, dev, world, rank, kw, workload_envs, args, value if world == 1 else None,
                                ms_total / K if world == 1 else None, loop, rollout)

    floor_us = launch_floor_us(torch, dev)
    moved = int(real_bytes_per_env_step(kw, s_o=s_o) * n_envs)     # what the resident layout really moves per launch
    copy_us, copy_bytes = copy_same_bytes_us(torch, dev, moved)
    # ---- e2e: host buffers through swarm_step_host (H2D actions + D2H obs/reward/done inside the timed region).
    # Headline format: uint8 observations (values 0 / 1 / U, pad 255) — the reference's obs are small integers (int64 on the
    # host, A2C/noVis_sim.py:168-170) that its callers cast to float32 themselves (A2C/train.py:21-24), so a byte per cell is
    # lossless and quarters the PCIe traffic that bounds this number.  The float32 form is measured and reported beside it.
    E = max(1, args.e2e_steps)
    a_h = torch.empty((n_act, n_envs, D), dtype=torch.uint8).pin_memory(); a_h.copy_(loop.acts.cpu())
    e8 = BatchedDroneEnv(n_envs, auto_reset=True, seed=2026, env_id_base=rank * args.sets * n_envs, obs_dtype=torch.uint8, **kw)
    e8_ms, e8_d2h = measure_e2e(torch, dist, dev, world, e8, a_h, torch.uint8, E)
    e8.close()
    ef = envs[0] if args.obs_dtype == "float32" else BatchedDroneEnv(n_envs, auto_reset=True, seed=2026, env_id_base=rank * args.sets * n_envs, **kw)
    ef_ms, ef_d2h = measure_e2e(torch, dist, dev, world, ef, a_h, torch.float32, E)
    if ef is not envs[0]:
        ef.close()
    # the packed 2-bit window format (SWARM_DTYPE_P2, 8 B per drone): reported beside the headline, never as it
    ep_ms = ep_d2h = None
    if kw.get("obs_mode") == "window" and kw.get("window") == 5 and D <= 32:
        try:
            ep = BatchedDroneEnv(n_envs, auto_reset=True, seed=2026, env_id_base=rank * args.sets * n_envs, obs_dtype="packed2", **kw)
            ep_ms, ep_d2h = measure_e2e(torch, dist, dev, world, ep, a_h, torch.int64, E)
            ep.close()
        except Exception as exc:                    # every rank takes the same branch: creation depends on the shape only
            print(f"[bench] packed2 e2e skipped: {exc!r}", file=sys.stderr)
    h2d = n_envs * D
    agent_e2e = float(n_envs) * D * E * world

    # ---- the statistics exchange, checked outside every timed region (N > 1): the mailboxes must hold the totals of ALL ranks
    stats_global, stats_check = None, "n/a (single rank)"
    if world > 1 and not use_nccl:
        torch.cuda.synchronize(dev)
        dist.barrier()
        tot = envs[0].stats.clone()
        dist.all_reduce(tot)                # checker only: NCCL sum of the local totals
        envs[0].push_stats()                # the in-step feed lags by one launch: publish the final totals explicitly
        torch.cuda.synchronize(dev)
        dist.barrier()
        try:
            stats_global = envs[0].global_stats(min_pushes=1).tolist()
            stats_check = "ok" if stats_global == tot.tolist() else f"MISMATCH: mailboxes {stats_global} != all-reduce {tot.tolist()}"
        except Exception as exc:
            stats_check = f"ERROR: {exc!r}"
        flag = torch.tensor([1 if stats_check == "ok" else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0 and stats_check == "ok":
            stats_check = "MISMATCH on another rank"
    elif world > 1:
        stats_check = "n/a (--stats-mode nccl)"
    else:
        st = envs[0].stats.tolist()
        stats_check = "ok" if st[1] == st[0] * D and st[0] % n_envs == 0 and st[0] > 0 else f"MISMATCH: {st}"

    if rank == 0:
        peak, peak_src = load_peak()
        bytes_per_launch = int(algorithmic_bytes_per_env_step(kw, s_o=s_o) * n_envs)
        achieved = bytes_per_launch / (kern_avg_ms * 1e-3) / 1e9
        e0 = envs[0]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "u1 coverage bits + anomaly list / f32 positions+obs (integer rewards)", "data": "synthetic",
            "config": {"workload": args.workload, "envs_per_gpu": n_envs, "total_envs": n_envs * world, **kw,
                       "obs_dtype": args.obs_dtype, "auto_reset": True, "reward_mode": "A2C", "reset_rng": "philox4x32-10",
                       "actions": "uint8 uniform{0..3} (16 draws + their opposites, zero net drift), pre-generated on device",
                       "l2": f"{args.sets} independent env batches rotated; working set {args.sets * (e0._obs.numel() * e0._obs.element_size() + n_envs * 700) / 1e6:.0f} MB > 126 MB L2",
                       "parallelism": f"env-sharded x{world}, no data-path collective; per-step episode statistics: " +
                                      ("published by the step kernel into every rank's mailbox over NVLink peer memory (posted stores at kernel start, no collective kernel)"
                                       if not use_nccl else f"swarm_stats_allreduce (NCCL) on a side stream every {max(args.stats_every, 1)} step(s)")},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": load_traffic(args.workload) if args.obs_dtype == "float32" else None, "peak_source": peak_src,
                         "kernel": "swarm::step_warp_kernel (swarm::step_cta_kernel for cfg5a/cfg5b)",
                         "kernel_ms_avg": kern_avg_ms, "kernel_ms_avg_how": "timed region / steps: the stream holds only step kernels back to back under graph replay",
                         "kernel_ms_event_pairs_avg": pair_avg_ms, "kernel_ms_event_pairs_median": kern_ms[len(kern_ms) // 2],
                         "kernel_ms_event_pairs_how": "separate EAGER pass, one CUDA-event pair per launch (diagnostic; includes Python launch gaps)",
                         "launch": loop.describe(K, plan) + (" (timed between event nodes inside the graph)" if timed is not None else ""),
                         "ms_per_step_1024_steps": ms_long / K_long,
                         "algorithmic_bytes_per_launch": bytes_per_launch,
                         "algorithmic_bytes_per_agent_step": algorithmic_bytes_per_env_step(kw, s_o=s_o) / D},
            "e2e": {"value": agent_e2e / (e8_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": e8_d2h,
                    "steps": E, "ms_per_step": e8_ms / E, "obs_format": "uint8 (lossless: the obs values are the integers 0 / 1 / U, pad 255)",
                    "api": "swarm_step_host (pinned host buffers: H2D actions, D2H obs + reward + done + statistics, stream sync per call)",
                    "d2h_gbs_per_rank": e8_d2h / (e8_ms / E * 1e-3) / 1e9},
            "e2e_float32_obs": {"value": agent_e2e / (ef_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": ef_d2h,
                                "steps": E, "ms_per_step": ef_ms / E, "obs_format": "float32 (what the reference's callers cast to)",
                                "d2h_gbs_per_rank": ef_d2h / (ef_ms / E * 1e-3) / 1e9},
            "e2e_packed2_obs": None if ep_ms is None else {
                "value": agent_e2e / (ep_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": ep_d2h, "steps": E,
                "ms_per_step": ep_ms / E, "d2h_gbs_per_rank": ep_d2h / (ep_ms / E * 1e-3) / 1e9,
                "obs_format": "packed 2-bit windows (SWARM_DTYPE_P2: one uint64 per drone = two bit planes of its 5x5 window, lossless; "
                              "swarm_b200.unpack_windows expands it) — an extra line for consumers that take the packed form, not the headline"},
            "rollout": rollout, "launch_floor_us": floor_us,
            "same_size_copy": {"us": copy_us, "bytes": copy_bytes, "gbs": copy_bytes / (copy_us * 1e-6) / 1e9,
                               "step_over_copy": (ms_total / K * 1e3) / copy_us,
                               "what": "a plain torch device copy of as many bytes (read + write) as one step really moves with this layout, "
                                       "graph-replayed like `value`: the memory system's answer for a launch of this size"},
            "strong_scaling": strong,
            "stats_check": stats_check, "stats_global_env_batch0": stats_global,
            "gpu_launches": launches, "clocks": dict(clocks, window=f"{hold * n_act} untimed steps + the {K} timed steps, continuous load"),
        }
        if world == 1 and not args.no_cpu_baseline:
            r = python_reference_throughput(kw, steps=2, warmup=1, budget_s=15.0)
            v, cores, sample, _ = cpu_port_throughput(kw, budget_s=8.0, scan_mode=0)
            port = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
            if r is not None:
                line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["procs"], "kind": "reference",
                                        "sample": reference_sample_text(r, args.workload)}
                line["cpu_baseline_port"] = port
            else:
                line["cpu_baseline"] = port
            v2, _, sample2, _ = cpu_port_throughput(kw, budget_s=3.0, scan_mode=1)
            line["cpu_baseline_windowed"] = {"value": v2, "unit": UNIT, "cores": cores, "kind": "port (ours: bounded-window scan, not the reference's algorithm)", "sample": sample2}
        emit(line)
    loop.close()
    teardown(dev, world, envs, comm)


if __name__ == "__main__":
    main()
