"""Freezes traces of the 2D simulator's semantic donors and of the vec-env worker protocol from the UNMODIFIED reference files
into tests/golden/ (rows a11 / a12 / a13 of SURVEY §8).  Run in the build container only:

    python tests/golden/make_donor_golden.py

Nothing that renders is touched: `swarmsim-2d/drone.py` and `goals.py` subclass `arcade.Sprite` only to be drawable, so a
three-line stand-in base class (no graphics, no arcade code) is enough to import them unchanged; `run_env.py` (window, sprite
collision) is NOT imported.  `A2C/envs.py` imports gym / baselines at module level for its Atari leftovers; empty stand-in
modules let `worker` (`envs.py:17-41`) be imported unchanged.

  drone_update_targets.npz  Drone.update (drone.py:29-59): per case the grid (-1 = mapped), the drone's position and the +-1
                            move it made on each axis; includes drones standing exactly on an unmapped cell (`if not min_dist`).
  goals_dwell.npz           NormalGoal / AnomalousGoal (goals.py:4-41) driven by the collision loop body of run_env.py:183-201:
                            per tick the (drone, cell) hits in drone order -> wait_steps of every cell and the mapped flags.
  vecenv_worker.npz         worker (envs.py:17-41) around the reference DroneEnv through a 4-tuple adapter: the stream of
                            ('reset'|'step', action) commands and the (ob, reward, done) replies, episodes ending inside the
                            trace (the reply to the finishing step carries the RESET observation, envs.py:24-26).
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def load(path, name, stubs):
    saved = {k: sys.modules.get(k) for k in stubs}
    sys.modules.update(stubs)
    try:
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod


def arcade_stub():
    m = types.ModuleType("arcade")

    class Sprite:                                   # stand-in base class: a drawable thing with a centre, nothing else
        def __init__(self, filename=None, scale=1):
            self.center_x, self.center_y = 0, 0
    m.Sprite = Sprite
    return {"arcade": m}


def drone_cases():
    drone_mod = load(os.path.join(REF, "swarmsim-2d", "drone.py"), "_ref_drone", arcade_stub())
    rs = np.random.RandomState(20261)
    MARGIN, W = 0, 1                                # goal_coords[row][col] = [col, row] + W//2 = lattice coordinates (drone.py:14-16)
    grids, poss, moves, rets, shapes = [], [], [], [], []
    for case in range(160):
        R, C = [(5, 5), (6, 9), (10, 10), (3, 4), (1, 7), (12, 5)][case % 6]
        frac = [0.0, 0.3, 0.6, 0.85, 0.97][case % 5]
        grid = np.where(rs.rand(R, C) < frac, -1, 1).astype(np.int64)
        if case % 7 == 3:
            grid[:] = -1                            # nothing left: update() returns 1 and does not move
        # positions on the half-integer lattice (what noVis step sizes 0.5 / 1.0 produce), some exactly on an unmapped cell
        x = rs.randint(0, 2 * R - 1) / 2.0
        y = rs.randint(0, 2 * C - 1) / 2.0
        if case % 3 == 0:
            free = np.argwhere(grid != -1)
            if len(free):
                x, y = (float(v) for v in free[rs.randint(len(free))])
        d = drone_mod.Drone(None, 1, grid.tolist(), 0, MARGIN, W, W, R, C)
        d.center_x, d.center_y = y, x               # pixel x <-> column (noVis y), pixel y <-> row (noVis x)
        ret = d.update()
        g = np.full((12, 12), -2, np.int64); g[:R, :C] = grid
        grids.append(g); poss.append([x, y]); shapes.append([R, C])
        moves.append([d.center_y - x, d.center_x - y])          # (dx, dy) in noVis coordinates
        rets.append(1 if ret else 0)
    np.savez_compressed(os.path.join(HERE, "donor", "drone_update_targets.npz"), grid=np.array(grids, np.int8), pos=np.array(poss),
                        move=np.array(moves), ret=np.array(rets, np.int8), shape=np.array(shapes, np.int32))
    print("drone_update_targets:", len(grids), "cases;", int(np.sum(rets)), "with nothing left;",
          int(np.sum([abs(m[0]) + abs(m[1]) == 2 for m in moves])), "diagonal moves")


def goal_cases():
    goals = load(os.path.join(REF, "swarmsim-2d", "goals.py"), "_ref_goals", arcade_stub())
    rs = np.random.RandomState(20262)
    out = {}
    for case, (R, C, D, dn, da, T) in enumerate([(4, 4, 3, 10, 35, 160), (3, 5, 6, 2, 5, 80), (5, 5, 12, 3, 7, 90), (2, 2, 4, 1, 0, 20)]):
        anom = rs.rand(R, C) < 0.3
        cells = [[goals.AnomalousGoal(None, 1, da) if anom[r][c] else goals.NormalGoal(None, 1, dn) for c in range(C)] for r in range(R)]
        mapped = np.zeros((R, C), bool)
        hits_all, wait_all, mapped_all = [], [], []
        for t in range(T):
            # the drones that touch a cell this tick, in drone order (run_env.py:169 `for drone in self.drone_sprite_list`)
            hits = np.full((D, 2), -1, np.int32)
            for i in range(D):
                if rs.rand() < 0.8:
                    hits[i] = [rs.randint(R), rs.randint(C)] if rs.rand() < 0.5 else [t % R, (t // 2 + i // 2) % C]
            for i in range(D):
                r, c = hits[i]
                if r < 0 or mapped[r][c]:
                    continue                        # removed sprites are no longer in the hit list (run_env.py:185,196)
                goal = cells[r][c]
                if goal.get_wait_steps() == 0:      # run_env.py:184-189 / 195-200
                    mapped[r][c] = True
                else:
                    goal.reduce_wait_steps()        # run_env.py:190-191 / 201-202
            hits_all.append(hits)
            wait_all.append([[cells[r][c].get_wait_steps() for c in range(C)] for r in range(R)])
            mapped_all.append(mapped.copy())
        out[f"anom_{case}"] = anom
        out[f"hits_{case}"] = np.array(hits_all)
        out[f"wait_{case}"] = np.array(wait_all, np.int32)
        out[f"mapped_{case}"] = np.array(mapped_all)
        out[f"cfg_{case}"] = np.array([R, C, D, dn, da, T], np.int32)
        print(f"goals_dwell case {case}: {int(mapped.sum())} of {R * C} cells mapped after {T} ticks")
    out["n_cases"] = np.int32(4)
    np.savez_compressed(os.path.join(HERE, "donor", "goals_dwell.npz"), **out)


class FourTuple:
    """gym-style adapter the worker expects: step -> (ob, reward, done, info) around the reference's 3-tuple DroneEnv."""

    def __init__(self, env):
        self.env = env

    def step(self, a):
        ob, r, d = self.env.step(a)
        return ob, r, d, {}

    def reset(self):
        return self.env.reset()


class FakePipe:
    def __init__(self, cmds):
        self.cmds, self.sent, self.closed = list(cmds), [], False

    def recv(self):
        return self.cmds.pop(0)

    def send(self, x):
        self.sent.append(x)

    def close(self):
        self.closed = True


def vecenv_cases():
    stubs = {"gym": types.ModuleType("gym"), "baselines": types.ModuleType("baselines"),
             "baselines.common": types.ModuleType("baselines.common"),
             "baselines.common.atari_wrappers": types.ModuleType("baselines.common.atari_wrappers"),
             "baselines.common.vec_env": types.ModuleType("baselines.common.vec_env")}
    stubs["gym"].spaces = types.ModuleType("gym.spaces")
    stubs["gym.spaces"] = stubs["gym"].spaces
    stubs["baselines.common.atari_wrappers"].make_atari = stubs["baselines.common.atari_wrappers"].wrap_deepmind = None
    stubs["baselines.common.vec_env"].VecEnv = object
    envs_mod = load(os.path.join(REF, "A2C", "envs.py"), "_ref_envs", stubs)
    nov = load(os.path.join(REF, "A2C", "noVis_sim.py"), "_ref_noVis_a2c", {})
    out = {}
    for case, (R, C, D, ss, T, seed) in enumerate([(3, 3, 3, 1.0, 160, 11), (4, 3, 2, 1.0, 200, 12), (3, 4, 5, 0.5, 700, 13)]):
        rs = np.random.RandomState(seed)
        acts = rs.randint(0, 4, (T, D))
        cmds = [("reset", None)] + [("step", list(a)) for a in acts] + [("close", None)]
        np.random.seed(1000 + seed)                 # the env's resets draw from the global stream (noVis_sim.py:48-49,60-61)
        states = []

        class Wrapper:
            @staticmethod
            def x():
                env = nov.DroneEnv(row_count=R, col_count=C, n_anamolous=2, n_drones=D, step_size=ss)
                ft = FourTuple(env)
                reset0 = ft.reset

                def reset():                        # record the state every reset produced (to inject on the device)
                    ob = reset0()
                    states.append((np.array(env.grid, np.int8).T.copy(), np.array(env.n_drones_pos, np.float64)))
                    return ob
                ft.reset = reset
                return ft
        remote, parent = FakePipe(cmds), FakePipe([])
        envs_mod.worker(remote, parent, Wrapper)    # the UNMODIFIED worker loop (envs.py:17-41), in-process
        assert remote.closed and parent.closed
        first, replies = remote.sent[0], remote.sent[1:]
        assert len(replies) == T
        out[f"cfg_{case}"] = np.array([R, C, D, T, seed], np.int32)
        out[f"ss_{case}"] = np.float64(ss)
        out[f"actions_{case}"] = acts.astype(np.uint8)
        out[f"obs0_{case}"] = np.asarray(first, np.int8)
        out[f"obs_{case}"] = np.array([np.asarray(r[0]) for r in replies], np.int8)
        out[f"rew_{case}"] = np.array([r[1] for r in replies], np.float64)
        out[f"done_{case}"] = np.array([r[2] for r in replies], np.uint8)
        out[f"reset_grid_{case}"] = np.array([s[0] for s in states], np.int8)
        out[f"reset_pos_{case}"] = np.array([s[1] for s in states], np.float64)
        print(f"vecenv_worker case {case}: {int(out[f'done_{case}'].sum())} episodes ended inside {T} steps, {len(states)} resets")
        assert out[f"done_{case}"].sum() >= 2 and len(states) == 1 + out[f"done_{case}"].sum()
    out["n_cases"] = np.int32(3)
    np.savez_compressed(os.path.join(HERE, "donor", "vecenv_worker.npz"), **out)


if __name__ == "__main__":
    drone_cases()
    goal_cases()
    vecenv_cases()
