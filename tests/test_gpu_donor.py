"""GPU: the CUDA path (through the C ABI) against the donor fixtures frozen from the UNMODIFIED reference files — the same
drivers as tests/test_donor_golden.py (rows a11 / a13 of SURVEY §8), with BatchedDroneEnv in place of the oracle; row a12's
target choice is pinned on the oracle and the CUDA greedy source is checked against that oracle here."""
import numpy as np
import pytest
import torch

import donor_common as dc

pytestmark = pytest.mark.gpu


class GpuEngine:
    def __init__(self, **kw):
        from swarm_b200 import BatchedDroneEnv
        self.g = BatchedDroneEnv(1, obs_dtype=torch.uint8, **kw)

    def set_state(self, grid, pos, t, wait):
        self.g.set_state(np.asarray(grid, np.int8), np.asarray(pos, np.float32), np.asarray(t, np.int32),
                         None if wait is None else np.asarray(wait, np.int16))

    def step(self, a):
        obs, rew, done, _ = self.g.step(torch.from_numpy(np.asarray(a, np.uint8)).to(self.g.device))
        return obs.reshape(1, -1).cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy()

    def state(self):
        g, p, t, w = self.g.get_state()
        return g.cpu().numpy(), p.cpu().numpy(), t.cpu().numpy(), None if w is None else w.cpu().numpy()


def test_dwell_counters_match_reference_goals_on_gpu():
    def factory(R, C, D, dn, da):
        return GpuEngine(row_count=R, col_count=C, n_drones=D, n_anamolous=0, step_size=1.0, dwell_normal=dn, dwell_anom=da, auto_reset=False)
    assert dc.check_dwell(factory) == 4


def test_vecenv_worker_protocol_matches_reference_on_gpu():
    def factory(R, C, D, ss):
        return GpuEngine(row_count=R, col_count=C, n_drones=D, n_anamolous=2, step_size=ss, auto_reset=True, seed=77)
    assert dc.check_worker_protocol(factory) >= 20


def test_greedy_source_uses_the_pinned_target_choice():
    """swarm_greedy_actions == oracle_greedy_actions (whose target choice is the pinned drone_update_target) on the fixture's
    own grids — including drones standing on an unmapped cell (`if not min_dist`, drone.py:39)."""
    from oracle.oracle import OracleBatch
    from swarm_b200 import BatchedDroneEnv
    z = dc.load("drone_update_targets.npz")
    for k in range(len(z["pos"])):
        R, C = (int(v) for v in z["shape"][k])
        grid_t = (z["grid"][k][:R, :C] != -1).astype(np.int8).T[None].copy()
        pos = z["pos"][k][None, None].astype(np.float64)
        o = OracleBatch(1, row_count=R, col_count=C, n_drones=1, n_anamolous=0, step_size=0.5)
        o.set_state(grid_t, pos, np.zeros(1, np.int32))
        g = BatchedDroneEnv(1, row_count=R, col_count=C, n_drones=1, n_anamolous=0, step_size=0.5, auto_reset=False)
        g.set_state(grid_t, pos.astype(np.float32), np.zeros(1, np.int32))
        assert g.greedy_actions().cpu().numpy().tolist() == o.greedy_actions().tolist(), k
        g.close()
