"""CPU: the reference arm of bench.py (the one leg that runs without a GPU) prints exactly ONE JSON line with the contract's keys;
the caller-row workloads state why their reference arm cannot travel; the product arm refuses to run without a GPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=400, env=None):
    return subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), *args], capture_output=True, text=True, timeout=timeout, cwd=REPO,
                          env=env)


@pytest.mark.timeout(500)
def test_reference_arm_line():
    # as under torchrun: OMP_NUM_THREADS=1 in the environment must NOT shrink the reference arm to one core
    p = _run("--impl", "reference", "--steps", "1", "--warmup", "1", env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "agent-steps/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1 and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["config"]["workload"] == "cfg3" and d["config"]["n_drones"] == 16 and d["config"]["row_count"] == 64
    cb, e2e = d["cpu_baseline"], d["e2e"]
    assert cb["kind"] in ("port", "reference") and cb["value"] == d["value"] and cb["sample"]
    assert cb["cores"] == len(os.sched_getaffinity(0))
    assert d["config"]["total_envs"] >= 1 and d["config"]["env_steps_per_step"] >= 1
    if os.path.exists(os.path.join(REPO, "baseline", "_ref", "A2C", "noVis_sim.py")):
        assert cb["kind"] == "reference" and "UNMODIFIED" in cb["sample"]
        assert d["cpu_port"]["kind"] == "port" and d["cpu_port"]["cores"] == cb["cores"] and d["cpu_port"]["value"] > d["value"]
    assert e2e["value"] == d["value"] and e2e["h2d_bytes_per_step"] == 0 and e2e["d2h_bytes_per_step"] == 0


@pytest.mark.timeout(900)
@pytest.mark.parametrize("workload", ["cfg4_a2c", "cfg4_dql"])
def test_caller_rows_reference_arm(workload):
    """The caller-row workloads' reference arm runs the UNMODIFIED reference loop from baseline/_ref on the host cores (or says why
    it cannot when the reference is not installed)."""
    p = _run("--impl", "reference", "--workload", workload, "--steps", "1", "--warmup", "0", timeout=800,
             env=dict(os.environ, SWARM_BENCH_CORES="2"))
    assert p.returncode == 0, p.stderr[-2000:]
    d = json.loads(p.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["config"]["workload"] == workload
    if os.path.exists(os.path.join(REPO, "baseline", "_ref", "A2C", "train.py")):
        assert d["value"] > 0 and d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] == 2
        assert "UNMODIFIED" in d["cpu_baseline"]["sample"] and d["e2e"]["value"] == d["value"]
    else:
        assert "unavailable" in d


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without a GPU")
@pytest.mark.parametrize("workload", ["cfg3", "cfg4_dql"])
def test_product_arm_has_no_cpu_fallback(workload):
    p = _run("--workload", workload, "--steps", "1", "--warmup", "1")
    assert p.returncode != 0 and "no CPU fallback" in (p.stderr + p.stdout)


def test_timed_region_is_graph_replay_at_any_step_count():
    """bench.py's timed region: floor(K / 32) replays of the 32-step graph + one replay of a (K mod 32)-step graph, and the
    `roofline.launch` label says what ran (--steps 20 was 20 eager launches labelled as a graph)."""
    sys.path.insert(0, REPO)
    import bench

    class FakeLoop(bench.StepLoop):
        def __init__(self):
            self.use_graph, self.graphs, self.made = True, {}, []

        def graph_for(self, n, tail=None):
            if n <= 0:
                return None
            self.made.append(n)
            return f"graph{n}"

    for K, want, label in [(20, (0, None, "graph20"), "CUDA graph replay: 1 x 20-step graph"),
                           (1024, (32, "graph32", None), "CUDA graph replay: 32 x 32-step graph"),
                           (70, (2, "graph32", "graph6"), "CUDA graph replay: 2 x 32-step graph + 1 x 6-step graph"),
                           (1, (0, None, "graph1"), "CUDA graph replay: 1 x 1-step graph")]:
        loop = FakeLoop()
        plan = loop.plan(K)
        assert plan == want
        assert loop.describe(K, plan) == label
    assert FakeLoop().describe(20, None).startswith("eager")
    src = open(os.path.join(REPO, "bench.py")).read()
    assert "os._exit" not in src.replace("no os._exit", "")


def test_byte_accounting_of_the_bench_line():
    """SURVEY §8(d) canonical bytes (the roofline numerator) and the bytes the round-2 layout really moves."""
    sys.path.insert(0, REPO)
    import bench
    kw3 = bench.WORKLOADS["cfg3"][1]
    assert bench.algorithmic_bytes_per_env_step(kw3) == 2325                      # 145.3 B per agent-step
    assert bench.real_bytes_per_env_step(kw3) == 512 + 32 + 256 + 32 + 16 + 5 + 1600   # plane, list, pos r+w, meta r+w, actions, reward+done, obs
    assert round(bench.real_bytes_per_env_step(kw3) * 65536 / 1e6, 1) == 160.8
    assert bench.algorithmic_bytes_per_env_step(bench.WORKLOADS["cfg2"][1]) == 5213
    assert bench.algorithmic_bytes_per_env_step(bench.WORKLOADS["cfg4"][1]) == 665
    assert bench.algorithmic_bytes_per_env_step(bench.WORKLOADS["cfg5a"][1]) == 328853
    for w in ("cfg2", "cfg4", "cfg5a"):                                           # 1 bit per cell: less than the canon's byte per cell
        kw = bench.WORKLOADS[w][1]
        assert bench.real_bytes_per_env_step(kw) <= bench.algorithmic_bytes_per_env_step(kw)
    kw = bench.WORKLOADS["cfg5b"][1]                                              # the whole 8 KiB plane is staged to serve 64 windows
    assert bench.real_bytes_per_env_step(kw) == 8192 + 32 + 1024 + 32 + 64 + 5 + 6400
