"""Freezes traces of the LIVE, UNMODIFIED reference env into tests/golden/*.npz (run in the build container only:
`python tests/golden/make_golden.py`; /root/reference does not exist on the GPU box).

Each trace: np.random.seed(seed); env = DroneEnv(**kw) from A2C/noVis_sim.py or DQL_model/noVis_sim.py; actions from
np.random.RandomState(1234 + seed).randint(0, 4, (T, D)); the env is stepped T times WITHOUT reset (so the
post-done behaviour of A2C/train.py:264 is covered) unless reset_on_done, in which case env.reset() follows a done
(A2C/train.py:78-80) and the reset state is recorded for injection.
Recorded per step: obs (int8, the reference's grid.T.flatten()), reward (f64), done, positions (f64), step_func_count.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refload  # noqa: E402

CASES = [
    # name, variant, seed, T, reset_on_done, kwargs
    ("cfg1_defaults_a2c", "A2C", 0, 700, False, {}),
    ("cfg1_defaults_dql", "DQL", 1, 700, False, {}),
    ("trainer_5x5_a2c", "A2C", 2, 400, True, dict(row_count=5, col_count=5, n_anamolous=5, n_drones=3, step_size=1.0)),
    ("trainer_5x5_dql", "DQL", 3, 400, True, dict(row_count=5, col_count=5, n_anamolous=5, n_drones=3, step_size=1.0)),
    ("rect_6x9_d4", "A2C", 4, 500, False, dict(row_count=6, col_count=9, n_anamolous=3, n_drones=4, step_size=1.0)),
    ("rect_9x4_d2_dql", "DQL", 5, 400, True, dict(row_count=9, col_count=4, n_anamolous=2, n_drones=2, step_size=0.5)),
    ("cd1p2_8x5_d9", "A2C", 6, 300, False, dict(row_count=8, col_count=5, n_anamolous=4, n_drones=9, step_size=0.5, collision_dist=1.2)),
    ("cd1p6_7x7_d2", "DQL", 7, 500, False, dict(row_count=7, col_count=7, n_anamolous=5, n_drones=2, step_size=0.25, collision_dist=1.6)),
    ("cd0p4_6x6_d5", "A2C", 8, 300, True, dict(row_count=6, col_count=6, n_anamolous=6, n_drones=5, step_size=1.0, collision_dist=0.4)),
    ("cfg4_10x10_d8", "A2C", 9, 600, True, dict(row_count=10, col_count=10, n_anamolous=5, n_drones=8, step_size=1.0)),
    ("cfg2_32x32_d4", "A2C", 10, 1500, False, dict(row_count=32, col_count=32, n_anamolous=5, n_drones=4, step_size=1.0)),
    ("d40_12x12", "A2C", 11, 200, True, dict(row_count=12, col_count=12, n_anamolous=7, n_drones=40, step_size=1.0)),
    ("single_drone_4x4", "DQL", 12, 200, True, dict(row_count=4, col_count=4, n_anamolous=1, n_drones=1, step_size=1.0)),
]
DEFAULTS = dict(row_count=10, col_count=10, n_anamolous=5, uncertainity=5, collision_dist=0.5, n_drones=3, step_size=0.5)


def make(name, variant, seed, T, reset_on_done, kw):
    cfg = dict(DEFAULTS, **kw)
    env = refload.make_ref_env(variant, seed=seed, **cfg)
    D, RC = cfg["n_drones"], cfg["row_count"] * cfg["col_count"]
    actions = np.random.RandomState(1234 + seed).randint(0, 4, (T, D)).astype(np.uint8)
    g0, p0, t0 = refload.export_state(env)
    obs = np.zeros((T, RC), np.int8)
    rew = np.zeros(T, np.float64)
    done = np.zeros(T, np.uint8)
    pos = np.zeros((T, D, 2), np.float64)
    tcount = np.zeros(T, np.int32)
    reset_steps, reset_grid, reset_pos = [], [], []
    for k in range(T):
        o, r, d = env.step([int(a) for a in actions[k]])
        obs[k], rew[k], done[k] = o, r, d
        pos[k] = np.array(env.n_drones_pos, np.float64)
        tcount[k] = env.step_func_count
        if d and reset_on_done:
            env.reset()
            g, p, _ = refload.export_state(env)
            reset_steps.append(k)
            reset_grid.append(g)
            reset_pos.append(p)
    out = os.path.join(HERE, name + ".npz")
    np.savez_compressed(
        out, config=json.dumps(dict(cfg, variant=variant, seed=seed, reset_on_done=reset_on_done)), grid0=g0, pos0=p0, t0=t0,
        actions=actions, obs=obs, reward=rew, done=done, pos=pos, t=tcount, reset_steps=np.array(reset_steps, np.int32),
        reset_grid=np.array(reset_grid, np.int8).reshape(len(reset_steps), cfg["col_count"], cfg["row_count"]),
        reset_pos=np.array(reset_pos, np.float64).reshape(len(reset_steps), D, 2))
    print(f"{name}: T={T} dones={int(done.sum())} resets={len(reset_steps)} reward_sum={rew.sum():.0f} -> {os.path.getsize(out)} B")


if __name__ == "__main__":
    assert refload.have_reference(), "needs /root/reference"
    for case in CASES:
        make(*case)
