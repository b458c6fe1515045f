"""Drivers shared by the CPU (oracle) and GPU (CUDA through the C ABI) checks against the donor fixtures frozen from the
unmodified reference by tests/golden/make_donor_golden.py:

  drone_update_targets.npz   swarmsim-2d/drone.py:29-59  (row a12)
  goals_dwell.npz            swarmsim-2d/goals.py:4-41 driven by the loop body of run_env.py:183-201  (row a13)
  vecenv_worker.npz          A2C/envs.py:17-41 worker around the reference DroneEnv  (row a11)

An "engine" is anything with the OracleBatch surface used here (the oracle itself, or tests' GpuEngine adapter around
BatchedDroneEnv): set_state(grid, pos, t, wait), step(actions) -> (obs, rew, done), state() -> (grid, pos, t, wait).
"""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "donor")
U = 5


def load(name):
    z = np.load(os.path.join(GOLDEN, name))
    return {k: z[k] for k in z.files}


# ---------------------------------------------------------------------------------------------------- a13: dwell goals
def dwell_cases():
    z = load("goals_dwell.npz")
    for c in range(int(z["n_cases"])):
        R, C, D, dn, da, T = (int(v) for v in z[f"cfg_{c}"])
        yield dict(R=R, C=C, D=D, dn=dn, da=da, T=T, anom=z[f"anom_{c}"], hits=z[f"hits_{c}"], wait=z[f"wait_{c}"], mapped=z[f"mapped_{c}"])


def dwell_initial_state(case):
    """grid_t int8 [1, C, R] (value U on the fixture's anomalous cells), wait int16 [1, C, R] = the goals' initial wait_steps."""
    grid = np.where(case["anom"], U, 1).astype(np.int8)             # [R, C] = reference grid[row][col]
    wait = np.where(case["anom"], case["da"], case["dn"]).astype(np.int16)
    return grid.T[None].copy(), wait.T[None].copy()


def dwell_tick_inputs(case, t):
    """Positions + actions that make drone i end this tick exactly on its fixture cell (or on no cell at all): the drone is
    placed one step beside the cell and moved onto it; a drone without a hit sits at a half-integer point (distance to every
    lattice cell >= 0.5, so `dist < collision_dist` never holds, A2C/noVis_sim.py:81)."""
    R, C, D = case["R"], case["C"], case["D"]
    pos = np.zeros((1, D, 2), np.float64)
    act = np.zeros((1, D), np.uint8)
    for i in range(D):
        r, c = (int(v) for v in case["hits"][t][i])
        if r < 0:
            pos[0, i] = [0.5, 0.5]; act[0, i] = 2 if C > 1 else 1    # stays on the half-integer lattice in at least one axis
            if C == 1 and R == 1:
                raise ValueError("1x1 grids cannot park a drone")
        elif c > 0:
            pos[0, i] = [r, c - 1]; act[0, i] = 2                     # y += 1 -> (r, c)
        elif C > 1:
            pos[0, i] = [r, c + 1]; act[0, i] = 0                     # y -= 1 -> (r, c)
        elif r > 0:
            pos[0, i] = [r - 1, c]; act[0, i] = 1                     # x += 1
        else:
            pos[0, i] = [r + 1, c]; act[0, i] = 3                     # x -= 1
    return pos, act


def check_dwell(engine_factory):
    """engine_factory(R, C, D, dwell_normal, dwell_anom) -> engine (one env, step_size 1.0, cd 0.5, auto_reset off)."""
    n = 0
    for case in dwell_cases():
        eng = engine_factory(case["R"], case["C"], case["D"], case["dn"], case["da"])
        grid, wait = dwell_initial_state(case)
        pos0 = np.zeros((1, case["D"], 2))
        eng.set_state(grid, pos0, np.zeros(1, np.int32), wait)
        for t in range(case["T"]):
            pos, act = dwell_tick_inputs(case, t)
            g, _, _, w = eng.state()
            eng.set_state(g, pos, np.full(1, t, np.int32), w)
            eng.step(act)
            g, _, _, w = eng.state()
            mapped = (np.asarray(g)[0].T == 0)
            assert np.array_equal(mapped, case["mapped"][t]), f"case {n} tick {t}: mapped cells differ from goals.py / run_env.py:183-201"
            live = ~mapped
            assert np.array_equal(np.asarray(w)[0].T[live], case["wait"][t][live]), f"case {n} tick {t}: wait_steps differ"
        n += 1
    return n


# ---------------------------------------------------------------------------------------------------- a11: vec-env worker
def worker_cases():
    z = load("vecenv_worker.npz")
    for c in range(int(z["n_cases"])):
        R, C, D, T, seed = (int(v) for v in z[f"cfg_{c}"])
        yield dict(R=R, C=C, D=D, T=T, ss=float(z[f"ss_{c}"]), actions=z[f"actions_{c}"], obs0=z[f"obs0_{c}"], obs=z[f"obs_{c}"],
                   rew=z[f"rew_{c}"], done=z[f"done_{c}"], reset_grid=z[f"reset_grid_{c}"], reset_pos=z[f"reset_pos_{c}"])


def check_worker_protocol(engine_factory):
    """engine_factory(R, C, D, step_size) -> engine with ONE env and auto_reset ON.  Follows the reference worker's command
    stream (envs.py:20-29): every reply must carry the reference's reward and done; a reply to a NON-finishing step carries the
    stepped observation; the reply to a FINISHING step carries the observation of the freshly reset env (envs.py:24-26) — the
    reference's reset state comes from numpy's MT19937, ours from Philox, so that observation is compared with OUR reset state
    and the reference's recorded reset state is then injected to keep following the trace."""
    n_done = 0
    for case in worker_cases():
        eng = engine_factory(case["R"], case["C"], case["D"], case["ss"])
        k_reset = 0
        eng.set_state(case["reset_grid"][0][None], case["reset_pos"][0][None], np.zeros(1, np.int32), None)
        g, _, _, _ = eng.state()
        assert np.array_equal(np.asarray(g)[0].reshape(-1), case["obs0"])          # reply to ('reset', None)
        for k in range(case["T"]):
            obs, rew, done = eng.step(case["actions"][k][None])
            assert float(rew[0]) == case["rew"][k], f"step {k}: reward {float(rew[0])} != reference {case['rew'][k]}"
            assert bool(done[0]) == bool(case["done"][k]), f"step {k}: done"
            g, _, t, _ = eng.state()
            if done[0]:
                # ours: the returned obs is the RESET obs (the env already holds its next episode, t == 0) ...
                assert int(np.asarray(t)[0]) == 0
                assert np.array_equal(np.asarray(obs)[0].astype(np.int8), np.asarray(g)[0].reshape(-1)), f"step {k}: obs is not the reset obs"
                # ... and so is the reference's: its reply equals the state its own reset produced
                k_reset += 1
                assert np.array_equal(case["obs"][k], case["reset_grid"][k_reset].reshape(-1))
                eng.set_state(case["reset_grid"][k_reset][None], case["reset_pos"][k_reset][None], np.zeros(1, np.int32), None)
                n_done += 1
            else:
                assert np.array_equal(np.asarray(obs)[0].astype(np.int8), case["obs"][k]), f"step {k}: obs"
    return n_done
