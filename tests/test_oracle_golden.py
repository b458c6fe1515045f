"""CPU: the oracle (oracle/swarm_oracle.c) against the traces frozen from the live reference (tests/golden/)."""
import numpy as np
import pytest

from common import env_kwargs, golden_names, load_golden
from oracle.oracle import OracleBatch


@pytest.mark.parametrize("scan_mode", [0, 1])
@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_golden(name, scan_mode):
    g = load_golden(name)
    kw, rm = env_kwargs(g["config"])
    o = OracleBatch(1, reward_mode=rm, scan_mode=scan_mode, **kw)
    o.set_state(g["grid0"][None], g["pos0"][None], [int(g["t0"])])
    resets = {int(s): i for i, s in enumerate(g["reset_steps"])}
    for k in range(g["actions"].shape[0]):
        obs, rew, done = o.step(g["actions"][k][None])
        assert np.array_equal(obs[0].astype(np.int8), g["obs"][k]), (name, k)
        assert float(rew[0]) == g["reward"][k], (name, k)          # exact: rewards are small integers
        assert int(done[0]) == int(g["done"][k]), (name, k)
        assert np.array_equal(o.pos[0], g["pos"][k]), (name, k)      # exact: dyadic step sizes
        assert int(o.meta[0, 0]) == int(g["t"][k])
        if k in resets:
            i = resets[k]
            o.set_state(g["reset_grid"][i][None], g["reset_pos"][i][None], [0])


def test_golden_covers_edge_cases():
    names = golden_names()
    assert len(names) >= 12
    assert any(load_golden(n)["done"].sum() > 50 for n in names), "stepping past done must be covered"
    assert any(len(load_golden(n)["reset_steps"]) > 0 for n in names), "reset after done must be covered"
