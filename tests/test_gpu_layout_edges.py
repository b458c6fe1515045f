"""GPU: edge cases of the round-2 state layout (1-bit coverage map + anomaly list / anomaly plane) and of the W = 5 window fast
paths, in lockstep with the oracle: list capacity (0 and 8 anomalies), PLANE mode (more than 8 anomalies, a side longer than 256),
rows that are / are not whole words, grids smaller than the window, every tile size, uint8 windows, the CTA kernel's window path
with drone counts that do / do not give a 16-byte obs block, and injected mid-episode states whose found anomalies read 0."""
import numpy as np
import pytest
import torch

from test_gpu_parity import make_pair, run_lockstep, assert_state_equal

pytestmark = pytest.mark.gpu

W5 = dict(obs_mode=1, window=5)


@pytest.mark.parametrize("kw", [
    dict(row_count=20, col_count=20, n_drones=8, step_size=1.0, **W5),                     # rows not whole words, TS = 8
    dict(row_count=10, col_count=10, n_drones=3, step_size=0.5, **W5),                     # TS = 4, half-integer positions
    dict(row_count=32, col_count=32, n_drones=4, step_size=1.0, **W5),                     # rows = one word
    dict(row_count=96, col_count=17, n_drones=16, step_size=1.0, **W5),                    # rows = three words, TS = 16
    dict(row_count=33, col_count=40, n_drones=32, step_size=1.0, **W5),                    # TS = 32
    dict(row_count=3, col_count=4, n_drones=5, step_size=1.0, n_anamolous=2, **W5),        # grid smaller than the window: pad on both sides
    dict(row_count=1, col_count=6, n_drones=2, step_size=1.0, n_anamolous=1, **W5),
    dict(row_count=64, col_count=64, n_drones=16, step_size=1.0, n_anamolous=8, **W5),     # full list
    dict(row_count=64, col_count=64, n_drones=16, step_size=1.0, n_anamolous=0, **W5),     # empty list
    dict(row_count=12, col_count=9, n_drones=6, step_size=1.0, n_anamolous=7),             # full obs, 7 listed cells patched
    dict(row_count=10, col_count=10, n_drones=3, step_size=1.0, n_anamolous=12, **W5),     # PLANE mode (> 8 anomalies): CTA kernel
    dict(row_count=10, col_count=10, n_drones=3, step_size=1.0, n_anamolous=30),           # PLANE mode, full obs
    dict(row_count=300, col_count=3, n_drones=4, step_size=1.0, n_anamolous=3, **W5),      # PLANE mode (side > 256)
    dict(row_count=260, col_count=2, n_drones=2, step_size=1.0, n_anamolous=2),
    dict(row_count=40, col_count=40, n_drones=64, step_size=1.0, **W5),                    # CTA kernel, W = 5 image path (D % 4 == 0)
    dict(row_count=40, col_count=40, n_drones=33, step_size=1.0, **W5),                    # CTA kernel, obs block not a 16 B multiple
    dict(row_count=256, col_count=256, n_drones=40, step_size=1.0, n_anamolous=8, **W5),   # largest LIST-mode sides
])
@pytest.mark.parametrize("auto_reset", [False, True])
def test_lockstep_layout_edges(kw, auto_reset):
    N = 61 if kw["row_count"] * kw["col_count"] < 10000 else 9
    g, o = make_pair(N, auto_reset=auto_reset, seed=41, **kw)
    run_lockstep(g, o, 120, seed=12, check_every=2)
    g.close()


def test_uint8_windows_every_tile_size():
    for D, R in ((3, 10), (8, 20), (16, 64), (32, 33)):
        g, o = make_pair(45, auto_reset=True, seed=5, obs_dtype=torch.uint8, row_count=R, col_count=R, n_drones=D, step_size=1.0, **W5)
        rs = np.random.RandomState(1)
        for k in range(80):
            a = rs.randint(0, 4, (45, D)).astype(np.uint8)
            obs, rew, done, _ = g.step(torch.from_numpy(a).cuda())
            oo, orr, od = o.step(a)
            want = np.where(oo < 0, 255, oo).astype(np.uint8)
            assert np.array_equal(obs.reshape(45, -1).cpu().numpy(), want), (D, k)
            assert np.array_equal(rew.cpu().numpy(), orr) and np.array_equal(done.cpu().numpy(), od)
        g.close()


@pytest.mark.parametrize("kw", [dict(row_count=16, col_count=16, n_drones=4, step_size=1.0, **W5),
                                dict(row_count=16, col_count=16, n_drones=4, step_size=1.0),
                                dict(row_count=24, col_count=24, n_drones=40, step_size=1.0, **W5)])
def test_injected_mid_episode_states(kw):
    """set_state with grids from the middle of an episode: some anomalous cells already found (they read 0, so the injected
    grid holds FEWER cells of value U than n_anamolous), others not; the list's spare entries must never turn a normal cell into
    an anomalous one.  Also the all-normal grid (no U at all) and the all-mapped-but-one grid."""
    N = 40
    g, o = make_pair(N, auto_reset=False, seed=3, **kw)
    R, C = kw["row_count"], kw["col_count"]
    rs = np.random.RandomState(7)
    grid = np.ones((N, C, R), np.int8)
    for n in range(N):
        k = n % 6                                            # 0 .. 5 cells of value U left
        cells = rs.choice(R * C, size=k + 40, replace=False)
        grid[n].reshape(-1)[cells[:k]] = 5
        grid[n].reshape(-1)[cells[k:]] = 0                   # 40 mapped cells (incl. the found anomalies)
    grid[1] = 1                                              # no U, nothing mapped: only an out-of-grid key is harmless
    grid[2] = 0; grid[2, 0, 0] = 1                           # one cell left
    pos = rs.randint(0, min(R, C), (N, kw["n_drones"], 2)).astype(np.float64)
    t = np.zeros(N, np.int32)
    g.set_state(grid, pos.astype(np.float32), t)
    o.set_state(grid, pos, t)
    assert g.device_error() == 0
    assert_state_equal(g, o, "after injection")
    run_lockstep(g, o, 150, seed=2, check_every=1)
    g.close()


def test_injected_state_with_too_many_anomalies_is_flagged():
    from swarm_b200 import BatchedDroneEnv
    g = BatchedDroneEnv(4, row_count=8, col_count=8, n_drones=2, n_anamolous=3, step_size=1.0, auto_reset=False)
    grid = np.ones((4, 8, 8), np.int8)
    grid[2, 0, :5] = 5                                       # five cells of value U, the config says three
    g.set_state(grid, np.zeros((4, 2, 2), np.float32), np.zeros(4, np.int32))
    assert g.device_error() & 8                              # SWARM_DEVERR_BAD_STATE
    g.close()
