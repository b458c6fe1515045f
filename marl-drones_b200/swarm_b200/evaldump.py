"""Evaluation dump in the reference's text format: A2C/a2c_eval.py:73-95,114-122 writes, per step, one line to
`drone_pos_a2c.txt` (comma-joined `str([x, y])` per drone), one to `state_a2c.txt` (comma-joined `str(row)` of the obs
reshaped to [G, G]) and once `uncertrain_cells_a2c.txt` (comma-joined `str((row, col))`); A2C/policy_vis.py:29-66 parses them.

The strings are what numpy 1.x / Python floats print (`[3.0, 4.0]`, `[1 1 5 0]`): numpy 2 would render the reference's
np.float64 list elements as `np.float64(3.0)`, which policy_vis.py's `float(line[0][1:])` cannot parse.
Rendering stays out of scope (no graphics stack on this box); the files are the hand-off to the reference's own viewer.
"""
import os

import numpy as np

DRONE_POS_FILE, STATE_FILE, UNCERTAIN_FILE = "drone_pos_a2c.txt", "state_a2c.txt", "uncertrain_cells_a2c.txt"


def format_drone_pos(pos):
    """pos: [D][2] -> '[x0, y0],[x1, y1],...'  (a2c_eval.py:77-79,115-117)."""
    return ",".join("[" + ", ".join(repr(float(v)) for v in p) + "]" for p in np.asarray(pos, dtype=np.float64).reshape(-1, 2))


def format_state(obs, grid_size):
    """obs: int [G*G] (the env's transposed-flat observation) -> '[a b c],[d e f],...'  (a2c_eval.py:84-87,119-122)."""
    rows = np.asarray(obs).astype(np.int64).reshape(grid_size, grid_size)
    return ",".join(str(r) for r in rows)


def format_uncertain(points):
    """points: iterable of (row, col) -> '(r0, c0),(r1, c1),...'  (a2c_eval.py:92-94)."""
    return ",".join(str((int(a), int(b))) for a, b in points)


class EvalDumpWriter:
    """Streams one env's trajectory into the three files of a2c_eval.py (created fresh in `directory`)."""

    def __init__(self, directory, grid_size):
        self.grid_size = grid_size
        os.makedirs(directory, exist_ok=True)
        self.paths = {k: os.path.join(directory, k) for k in (DRONE_POS_FILE, STATE_FILE, UNCERTAIN_FILE)}
        for p in self.paths.values():
            if os.path.exists(p):
                os.remove(p)                                            # a2c_eval.py:74-75,81-82,89-90
        self._pos = open(self.paths[DRONE_POS_FILE], "a")
        self._state = open(self.paths[STATE_FILE], "a")

    def write_uncertain(self, points):
        with open(self.paths[UNCERTAIN_FILE], "a") as f:
            f.write(format_uncertain(points) + "\n")

    def write_step(self, pos, obs):
        self._pos.write(format_drone_pos(pos) + "\n")
        self._state.write(format_state(obs, self.grid_size) + "\n")

    def close(self):
        self._pos.close()
        self._state.close()


def dump_episode(env, action_fn, directory, max_steps=100000):
    """a2c_eval.py:63-131 on the single-env facade (swarm_b200.DroneEnv): reset, log, then step `action_fn(obs, env)` until
    done, logging positions and state after every step.  Returns the number of steps."""
    g = env.row_count
    assert env.row_count == env.col_count, "the reference's eval format assumes a square grid (args.grid_size)"
    w = EvalDumpWriter(directory, g)
    obs = env.reset()
    w.write_step(env.n_drones_pos, obs)
    w.write_uncertain(env.uncertain_points.keys())
    done, step = False, 0
    while not done and step < max_steps:
        obs, _, done = env.step(action_fn(obs, env))
        w.write_step(env.n_drones_pos, obs)
        step += 1
    w.close()
    return step


def parse_dump(directory, grid_size):
    """The reading side, as A2C/policy_vis.py:29-66 does it: (positions [T][D][2] float, states [T][G][G] int, cells [(r, c)])."""
    pos = []
    for line in open(os.path.join(directory, DRONE_POS_FILE)).read().split("\n")[:-1]:
        vals = [float(tok.strip(" []")) for tok in line.split(",")]
        pos.append(np.array(vals).reshape(-1, 2))
    states = []
    for line in open(os.path.join(directory, STATE_FILE)).read().split("\n")[:-1]:
        vals = [int(tok.replace("[", "").replace("]", "")) for part in line.split(",") for tok in part.split()]
        states.append(np.array(vals).reshape(grid_size, grid_size))
    cells_txt = open(os.path.join(directory, UNCERTAIN_FILE)).read().strip()
    nums = [int(t) for t in cells_txt.replace("(", " ").replace(")", " ").replace(",", " ").split()]
    cells = list(zip(nums[0::2], nums[1::2]))
    return np.array(pos), np.array(states), cells
