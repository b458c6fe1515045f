"""GPU: seeded random configurations (shape, drones, step size, collision distance, obs mode / window / dtype, anomalies incl. PLANE
mode, reward variant, auto-reset, dwell, action dtype, env count) in lock-step against the oracle: obs, reward, done every step,
full state and statistics at the end.  The configurations are drawn once from a fixed seed, so a failure names its case."""
import os

import numpy as np
import pytest
import torch

from test_gpu_parity import assert_state_equal, make_pair

pytestmark = pytest.mark.gpu


def draw_cases(n=160, seed=2026):
    rs = np.random.RandomState(seed)
    pick = lambda xs: xs[rs.randint(len(xs))]
    sides = [1, 2, 3, 4, 5, 6, 7, 9, 10, 12, 16, 31, 32, 33, 40, 64]
    cases = []
    for i in range(n):
        R, Cc = pick(sides), pick(sides)
        if R * Cc < 2:
            Cc = 3
        kw = dict(row_count=R, col_count=Cc, n_drones=pick([1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 32, 33, 48, 70]),
                  step_size=pick([0.5, 1.0, 1.0, 0.25, 2.0]), collision_dist=pick([0.3, 0.5, 0.5, 0.5, 0.75, 1.2]),
                  obs_mode=pick([0, 1]), window=pick([1, 3, 5, 5, 7]), n_anamolous=min(pick([0, 1, 3, 5, 8, 9, 12]), R * Cc),
                  uncertainity=pick([2, 5, 9, 127]), reward_mode=pick([0, 1]), auto_reset=bool(pick([0, 1])),
                  dwell=pick([(0, 0), (0, 0), (0, 0), (2, 3), (1, 0), (3, 1)]), obs_dtype=pick([torch.float32, torch.uint8]))
        cases.append((i, pick([1, 3, 33, 100]), pick([torch.uint8, torch.int32, torch.int64]), kw))
    return cases


# SWARM_FUZZ_N / SWARM_FUZZ_SEED widen the hunt for one-off runs (e.g. SWARM_FUZZ_N=2000 SWARM_FUZZ_SEED=7 pytest tests/test_gpu_fuzz.py)
CASES = draw_cases(int(os.environ.get("SWARM_FUZZ_N", 160)), int(os.environ.get("SWARM_FUZZ_SEED", 2026)))


@pytest.mark.parametrize("case", CASES, ids=[f"case{c[0]}" for c in CASES])
def test_fuzz_lockstep(case):
    i, N, adt, kw = case
    kw = dict(kw)
    g, o = make_pair(N, seed=100 + i, **kw)
    u8 = kw["obs_dtype"] == torch.uint8
    rs = np.random.RandomState(i)
    D = o.D
    for k in range(48):
        a = rs.randint(0, 4, (N, D))
        obs, rew, done, _ = g.step(torch.from_numpy(a).cuda().to(adt))
        oo, orr, od = o.step(a.astype(np.uint8))
        want = oo.astype(np.int16).astype(np.uint8) if u8 else oo.astype(np.float32)      # pad -1 -> 255 for uint8
        assert np.array_equal(obs.reshape(N, -1).cpu().numpy(), want), (kw, N, adt, k)
        assert np.array_equal(rew.cpu().numpy(), orr.astype(np.float32)) and np.array_equal(done.cpu().numpy(), od), (kw, N, adt, k)
    assert_state_equal(g, o, str(kw))
    assert np.array_equal(g.stats.cpu().numpy(), o.stats), kw
    assert g.device_error() == 0
    g.close()


ROLLOUT_CASES = draw_cases(max(48, int(os.environ.get("SWARM_FUZZ_N", 160)) // 4), int(os.environ.get("SWARM_FUZZ_SEED", 2026)) + 51)


@pytest.mark.parametrize("case", ROLLOUT_CASES, ids=[f"case{c[0]}" for c in ROLLOUT_CASES])
def test_fuzz_rollout(case):
    """swarm_rollout(T) launches (T steps in ONE launch, state carried on chip) on the same kind of random configurations:
    every step's obs (or only the last one's), rewards and dones against T oracle steps, three launches in a row."""
    i, N, _, kw = case
    kw = dict(kw)
    g, o = make_pair(N, seed=500 + i, **kw)
    odt = kw["obs_dtype"]
    u8 = odt == torch.uint8
    rs = np.random.RandomState(1000 + i)
    D = o.D
    per = int(np.prod(g.obs_shape))
    for launch in range(3):
        T = int(rs.randint(2, 10))
        last_only = bool(rs.randint(2))
        a = rs.randint(0, 4, (T, N, D)).astype(np.uint8)
        obs_out = torch.empty(((1 if last_only else T),) + tuple(g.obs_shape), dtype=odt, device="cuda")
        _, rew, done = g.rollout(torch.from_numpy(a).cuda(), obs_out=obs_out, obs_last_only=last_only)
        for t in range(T):
            oo, orr, od = o.step(a[t])
            want = oo.astype(np.int16).astype(np.uint8) if u8 else oo.astype(np.float32)
            if not last_only or t == T - 1:
                got = obs_out[0 if last_only else t].reshape(N, -1).cpu().numpy()
                assert np.array_equal(got, want), (kw, N, launch, t)
            assert np.array_equal(rew[t].cpu().numpy(), orr.astype(np.float32)) and np.array_equal(done[t].cpu().numpy(), od), (kw, N, launch, t)
    assert_state_equal(g, o, str(kw))
    assert np.array_equal(g.stats.cpu().numpy(), o.stats), kw
    assert g.device_error() == 0
    g.close()


GREEDY_CASES = draw_cases(n=64, seed=4242)


@pytest.mark.parametrize("case", GREEDY_CASES, ids=[f"case{c[0]}" for c in GREEDY_CASES])
def test_fuzz_greedy_actions(case):
    """swarm_greedy_actions (row-outward word search) against the oracle's literal R*C scan of drone.py on random shapes: fresh
    envs (every drone stands on an unmapped cell: the `if not min_dist` quirk), fractional positions, nearly and fully mapped grids;
    greedy and random moves alternate so that the states vary."""
    i, N, adt, kw = case
    kw = dict(kw)
    kw["auto_reset"] = bool(i % 2)
    g, o = make_pair(N, seed=900 + i, **kw)
    rs = np.random.RandomState(i)
    D = o.D
    steps = 30 + 4 * min(kw["row_count"] * kw["col_count"], 60)
    for k in range(steps):
        a = g.greedy_actions()
        assert np.array_equal(a.cpu().numpy(), o.greedy_actions()), (kw, N, k)
        if k % 5 == 4:
            a = torch.from_numpy(rs.randint(0, 4, (N, D)).astype(np.uint8)).cuda()
        g.step(a, want_obs=False)
        o.step(a.cpu().numpy())
    assert_state_equal(g, o, str(kw))
    g.close()
