"""Eval dump format (swarm_b200.evaldump) against the reference's own writer expressions (A2C/a2c_eval.py:73-95,114-122) and
its parser (A2C/policy_vis.py:29-66)."""
import os
import re

import numpy as np
import pytest

from swarm_b200 import evaldump as ed
import refload

# What a2c_eval.py wrote under numpy 1.x for a 3x3 grid, two drones (hand-derived from the expressions at :77-94):
KAT_POS = "[0.0, 2.0],[1.0, 1.0]"
KAT_STATE = "[1 1 5],[0 1 1],[1 5 1]"
KAT_CELLS = "(2, 0),(1, 2)"


def test_known_answer_strings():
    assert ed.format_drone_pos([[0.0, 2.0], [1.0, 1.0]]) == KAT_POS
    assert ed.format_state(np.array([1, 1, 5, 0, 1, 1, 1, 5, 1]), 3) == KAT_STATE
    assert ed.format_uncertain([(2, 0), (1, 2)]) == KAT_CELLS


@pytest.mark.skipif(not refload.have_reference(), reason="live reference only in the build container")
def test_strings_equal_reference_expressions():
    env = refload.make_ref_env("A2C", seed=4, row_count=6, col_count=6, n_drones=3, step_size=1.0)
    obs = env.reset()
    rs = np.random.RandomState(0)
    for _ in range(20):
        # the reference's expressions, verbatim semantics (a2c_eval.py:77-79,84-87,92-94,115-122)
        want_pos = ",".join(list(map(lambda x: str(x), env.n_drones_pos.copy())))
        want_pos = re.sub(r"np\.float64\(([^)]*)\)", r"\1", want_pos)          # numpy 2 repr -> what numpy 1.x printed
        want_state = ",".join(list(map(lambda x: str(x), obs.copy().reshape(6, 6))))
        want_cells = ",".join(list(map(lambda x: str(x), list(env.uncertain_points.keys()))))
        assert ed.format_drone_pos(env.n_drones_pos) == want_pos
        assert ed.format_state(obs, 6) == want_state
        assert ed.format_uncertain(env.uncertain_points.keys()) == want_cells
        obs, _, _ = env.step([int(a) for a in rs.randint(0, 4, 3)])


def test_writer_parser_round_trip(tmp_path):
    w = ed.EvalDumpWriter(str(tmp_path), 4)
    rs = np.random.RandomState(1)
    pos = rs.randint(0, 4, (7, 3, 2)).astype(np.float64)
    states = rs.choice([0, 1, 5], size=(7, 16))
    w.write_uncertain([(0, 1), (3, 3)])
    for p, s in zip(pos, states):
        w.write_step(p, s)
    w.close()
    gp, gs, cells = ed.parse_dump(str(tmp_path), 4)
    assert np.array_equal(gp, pos) and np.array_equal(gs, states.reshape(7, 4, 4)) and cells == [(0, 1), (3, 3)]


@pytest.mark.gpu
def test_dump_greedy_episode_on_facade(tmp_path):
    """A scripted greedy episode (swarm_greedy_actions, the non-learned baseline of SURVEY §8f-4) dumped from the CUDA facade:
    it terminates, the last state is fully mapped, and the files parse with the reference's reader."""
    from swarm_b200 import DroneEnv
    np.random.seed(11)
    env = DroneEnv(row_count=5, col_count=5, n_drones=3, step_size=1.0)
    steps = ed.dump_episode(env, lambda obs, e: e._env.greedy_actions()[0].tolist(), str(tmp_path), max_steps=500)
    pos, states, cells = ed.parse_dump(str(tmp_path), 5)
    assert 0 < steps < 500 and len(pos) == steps + 1 and len(states) == steps + 1
    assert states[-1].max() == 0 and len(cells) == 5
    assert set(np.unique(states[0])) <= {0, 1, 5}
