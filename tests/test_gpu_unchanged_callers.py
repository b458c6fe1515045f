"""Row N1 of VERDICT r1 / north_star: the reference's OWN callers run UNCHANGED on the CUDA env.

`baseline/_ref/` holds verbatim copies of /root/reference/{A2C,DQL_model} (installed by __graft_entry__.build(), git-ignored,
shipped to the GPU box with the snapshot).  Each test executes an unmodified reference script / function twice in fresh
interpreters with the same np.random / random / torch seeds — once on the reference's own CPU `noVis_sim.DroneEnv`, once with
`sys.modules['noVis_sim']` pre-seeded with the CUDA facade (swarm_b200.launcher) — and requires the two runs to agree:

  * DQL_model/DQLearning.py (`:147-190`, the module-level episode loop; `env.step(action)` `:170`): the printed
    `episode/timestep/reward/e` lines must be identical, line for line, up to the end of the episode;
  * A2C/train.py::train (`:35-157`; `env.reset()` `:53,80`, `env.step(actions)` `:76` with a list of ndarray(1,)): the
    (obs, reward, done) stream, the positions after every step and the logged loss / mean-reward lines must be identical.

Identical streams are the strongest possible statement: every action the nets pick depends on every observation returned
before it, so one wrong cell anywhere diverges the whole run.
"""
import json
import os
import subprocess
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(REPO, "baseline", "_ref")
PKG = os.path.join(REPO, "marl-drones_b200")

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(os.path.join(REF, "A2C", "train.py")),
                                 reason="baseline/_ref is not installed (run __graft_entry__.build() where /root/reference exists)")]

_DQL_DRIVER = r"""
import random, runpy, sys
import numpy as np, torch
seed, variant, script = int(sys.argv[1]), sys.argv[2], sys.argv[3]
np.random.seed(seed); random.seed(seed); torch.manual_seed(seed)
torch.set_num_threads(1)
if variant == "cuda":
    sys.path.insert(0, {pkg!r})
    from swarm_b200 import launcher
    launcher.main(["--variant", "DQL", script] + sys.argv[4:])          # the unmodified script on the CUDA facade
    import swarm_b200
    print("LOADED", swarm_b200._lib.SO_PATH, file=sys.stderr)
else:
    sys.argv = [script] + sys.argv[4:]
    sys.path.insert(0, {refdir!r})
    runpy.run_path(script, run_name="__main__")                          # the unmodified script on its own CPU env
"""


def _run_dql(variant, seed, extra, timeout=900):
    script = os.path.join(REF, "DQL_model", "DQLearning.py")
    code = _DQL_DRIVER.format(pkg=PKG, refdir=os.path.join(REF, "DQL_model"))
    p = subprocess.run([sys.executable, "-c", code, str(seed), variant, script] + extra, capture_output=True, text=True,
                       timeout=timeout, cwd=REPO)
    assert p.returncode == 0, p.stderr[-3000:]
    return [l for l in p.stdout.splitlines() if l.startswith("episode:")], p.stderr


@pytest.mark.timeout(1800)
@pytest.mark.parametrize("seed,grid", [(0, 3), (4, 3), (3, 5)])
def test_dqlearning_script_runs_unchanged_and_matches_cpu_reference(seed, grid):
    extra = ["--episodes", "1", "--grid_size", str(grid), "--n_drones", "3", "--n_anamolous", "2"]
    ref_lines, _ = _run_dql("ref", seed, extra)
    cuda_lines, err = _run_dql("cuda", seed, extra)
    assert "libswarm_b200.so" in err
    assert len(ref_lines) >= 1 and ref_lines[-1].count("reward: 1000") == 1      # the episode ran to `done` (DQLearning.py:171,186)
    assert cuda_lines == ref_lines


_A2C_DRIVER = r"""
import json, logging, os, sys, tempfile, types
import numpy as np, torch
seed, variant, grid, n_drones, total, T, policy, icm = int(sys.argv[1]), sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]), sys.argv[7], int(sys.argv[8])
refdir = {refdir!r}
if variant == "cuda":
    sys.path.insert(0, {pkg!r})
    from swarm_b200 import launcher
    mod = launcher.install("A2C")                    # sys.modules['noVis_sim'] = the CUDA facade module
    from noVis_sim import DroneEnv                   # what A2C/main.py:5 does
    assert DroneEnv.__module__.startswith("swarm_b200"), DroneEnv.__module__
sys.path.insert(1, refdir)
if variant != "cuda":
    from noVis_sim import DroneEnv                   # the reference's own CPU env (A2C/noVis_sim.py)
    assert DroneEnv.__module__ == "noVis_sim"
import models, train                                  # UNMODIFIED A2C/models.py, A2C/train.py
torch.set_num_threads(1)
os.chdir(tempfile.mkdtemp())                          # train() writes A2C_<policy>.log into the cwd
np.random.seed(seed); torch.manual_seed(seed)
env = DroneEnv(row_count=grid, col_count=grid, step_size=1.0, n_anamolous=5, n_drones=n_drones)    # A2C/main.py:97-103
nets = [models.Policy(env.state_size, n_drones, env.action_size, policy_type=policy) for _ in range(n_drones)]
opts = [torch.optim.AdamW(nets[i].parameters(), lr=2e-4) for i in range(n_drones)]
rec = []
step0, reset0 = env.step, env.reset
def step(actions):
    o, r, d = step0(actions)
    rec.append(["step", [int(np.asarray(a).reshape(-1)[0]) for a in actions], np.asarray(o).tolist(), float(r), bool(d),
                np.array(env.n_drones_pos, dtype=np.float64).tolist()])
    return o, r, d
def reset():
    o = reset0()
    rec.append(["reset", np.asarray(o).tolist(), np.array(env.n_drones_pos, dtype=np.float64).tolist(), str(np.asarray(o).dtype)])
    return o
env.step, env.reset = step, reset                    # instance attributes: the env class and train() stay untouched
args = types.SimpleNamespace(enable_icm=bool(icm), policy=policy, total_steps=total, rollout_steps=T, gamma=0.99, value_coeff=0.5,
                             entropy_coeff=0.01, grad_acc=1, grad_norm_limit=10.0, save_freq=10 ** 9 + 7, grid_size=grid)
train.train(args, nets, opts, env, env.state_size, n_drones)
logging.shutdown()
log = open([f for f in os.listdir(".") if f.endswith(".log")][0]).read().strip().splitlines()
digest = [float(sum(p.detach().double().abs().sum() for p in net.parameters())) for net in nets]
print("RESULT " + json.dumps(dict(rec=rec, log=log, digest=digest)))
if variant == "cuda":
    import swarm_b200
    print("LOADED", swarm_b200._lib.SO_PATH, file=sys.stderr)
"""


def _run_a2c(variant, seed, grid, n_drones, total, T, policy="MLP", icm=0, timeout=900):
    code = _A2C_DRIVER.format(pkg=PKG, refdir=os.path.join(REF, "A2C"))
    p = subprocess.run([sys.executable, "-c", code, str(seed), variant, str(grid), str(n_drones), str(total), str(T), policy, str(icm)],
                       capture_output=True, text=True, timeout=timeout, cwd=REPO)
    assert p.returncode == 0, p.stderr[-3000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):]), p.stderr


@pytest.mark.timeout(1800)
@pytest.mark.parametrize("seed,grid,n_drones,total,T,policy,icm", [
    (3, 5, 3, 100, 25, "MLP", 0),        # A2C/main.py defaults: 5x5, 3 drones, rollout 25 -> 4 outer iterations with updates
    (5, 5, 2, 60, 20, "MLP", 1),         # with the reference's ICM bonus on (train.py:85-92)
    (7, 5, 3, 40, 20, "CNN", 0),         # the CNN policy (hard-wired 5x5, models.py:120-160), host PyTorch on the facade
])
def test_train_runs_unchanged_and_matches_cpu_reference(seed, grid, n_drones, total, T, policy, icm):
    ref, _ = _run_a2c("ref", seed, grid, n_drones, total, T, policy, icm)
    got, err = _run_a2c("cuda", seed, grid, n_drones, total, T, policy, icm)
    assert "libswarm_b200.so" in err
    assert sum(1 for r in ref["rec"] if r[0] == "step") == total
    assert len(got["rec"]) == len(ref["rec"])
    for k, (a, b) in enumerate(zip(got["rec"], ref["rec"])):
        assert a == b, f"event {k} differs: {a[:2]} vs {b[:2]}"
    assert got["log"] == ref["log"]                  # loss / mean reward of every outer iteration, as logged by train.py:155
    assert got["digest"] == ref["digest"]            # weights after the AdamW updates
