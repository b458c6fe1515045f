"""CPU: the reference arm of bench.py (the one leg that runs without a GPU) prints exactly ONE JSON line with the contract's keys;
the caller-row workloads state why their reference arm cannot travel; the product arm refuses to run without a GPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=240):
    return subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), *args], capture_output=True, text=True, timeout=timeout, cwd=REPO)


@pytest.mark.timeout(300)
def test_reference_arm_line():
    p = _run("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "agent-steps/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1 and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["config"]["workload"] == "cfg3" and d["config"]["n_drones"] == 16 and d["config"]["row_count"] == 64
    cb, e2e = d["cpu_baseline"], d["e2e"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert e2e["value"] == d["value"] and e2e["h2d_bytes_per_step"] == 0 and e2e["d2h_bytes_per_step"] == 0


@pytest.mark.parametrize("workload", ["cfg4_a2c", "cfg4_dql"])
def test_caller_rows_reference_arm_says_unavailable(workload):
    p = _run("--impl", "reference", "--workload", workload)
    assert p.returncode == 0
    d = json.loads(p.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d and d["config"]["workload"] == workload


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without a GPU")
@pytest.mark.parametrize("workload", ["cfg3", "cfg4_dql"])
def test_product_arm_has_no_cpu_fallback(workload):
    p = _run("--workload", workload, "--steps", "1", "--warmup", "1")
    assert p.returncode != 0 and "no CPU fallback" in (p.stderr + p.stdout)
