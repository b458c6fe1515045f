"""CPU: the oracle against the donor fixtures frozen from the UNMODIFIED reference files (tests/golden/make_donor_golden.py):
rows a11 (vec-env worker protocol, A2C/envs.py:17-41), a12 (Drone.update target choice, swarmsim-2d/drone.py:29-59) and a13
(dwell counters, swarmsim-2d/goals.py:4-41 + run_env.py:183-201) of SURVEY §8."""
import numpy as np

import donor_common as dc
from oracle.oracle import OracleBatch


class OracleEngine:
    def __init__(self, **kw):
        self.o = OracleBatch(1, **kw)

    def set_state(self, grid, pos, t, wait):
        self.o.set_state(np.asarray(grid, np.int8), np.asarray(pos, np.float64), np.asarray(t, np.int32), wait)

    def step(self, a):
        return self.o.step(np.asarray(a, np.uint8))

    def state(self):
        return self.o.grid.copy(), self.o.pos.copy(), self.o.meta[:, 0].copy(), None if self.o.wait is None else self.o.wait.copy()


def test_drone_update_target_choice_matches_reference():
    z = dc.load("drone_update_targets.npz")
    n_quirk = 0
    for k in range(len(z["pos"])):
        R, C = (int(v) for v in z["shape"][k])
        ref_grid = z["grid"][k][:R, :C]                                  # -1 = mapped (drone.py:34)
        o = OracleBatch(1, row_count=R, col_count=C, n_drones=1, n_anamolous=0, step_size=0.5)
        o.grid[0] = (ref_grid != -1).astype(np.int8).T
        o.pos[0, 0] = z["pos"][k]
        target, direction = o.drone_update()
        assert (target[0, 0] < 0) == bool(z["ret"][k]), k                # `return 1` when nothing is left (drone.py:49-50)
        assert direction[0, 0].tolist() == z["move"][k].tolist(), (k, direction[0, 0], z["move"][k])
        x, y = z["pos"][k]
        if x == int(x) and y == int(y) and ref_grid[int(x), int(y)] != -1 and (ref_grid != -1).sum() > 1:
            n_quirk += 1                                                 # standing on an unmapped cell with others left
    assert n_quirk >= 10


def test_dwell_counters_match_reference_goals():
    def factory(R, C, D, dn, da):
        return OracleEngine(row_count=R, col_count=C, n_drones=D, n_anamolous=0, step_size=1.0, dwell_normal=dn, dwell_anom=max(da, 0) or 0,
                            auto_reset=False) if (dn or da) else OracleEngine(row_count=R, col_count=C, n_drones=D, n_anamolous=0, step_size=1.0)
    assert dc.check_dwell(factory) == 4


def test_vecenv_worker_protocol_matches_reference():
    def factory(R, C, D, ss):
        return OracleEngine(row_count=R, col_count=C, n_drones=D, n_anamolous=2, step_size=ss, auto_reset=True, seed=77)
    assert dc.check_worker_protocol(factory) >= 20
