"""CPU, build container only: the oracle against the LIVE unmodified reference (skipped where /root/reference is absent;
tests/golden/ carries the frozen equivalent to the GPU box)."""
import numpy as np
import pytest

import refload
from oracle.oracle import OracleBatch

pytestmark = pytest.mark.skipif(not refload.have_reference(), reason="live reference not present")

CASES = [
    (10, 10, 3, 0.5, 0.5, 5), (5, 5, 3, 0.5, 1.0, 5), (6, 9, 4, 0.4, 1.0, 3), (8, 5, 9, 1.2, 0.5, 4),
    (7, 7, 2, 1.6, 0.25, 5), (12, 12, 5, 0.5, 1.0, 7), (16, 16, 4, 0.5, 1.0, 5), (3, 3, 1, 0.5, 1.0, 9),
]


@pytest.mark.parametrize("variant,rm", [("A2C", 0), ("DQL", 1)])
@pytest.mark.parametrize("R,Cc,D,cd,ss,na", CASES)
def test_step_parity_live(variant, rm, R, Cc, D, cd, ss, na):
    seed = 77 + R * 13 + D
    kw = dict(row_count=R, col_count=Cc, n_anamolous=na, uncertainity=5, collision_dist=cd, n_drones=D, step_size=ss)
    env = refload.make_ref_env(variant, seed=seed, **kw)
    oracles = [OracleBatch(1, reward_mode=rm, scan_mode=s, **kw) for s in (0, 1)]
    g, p, t = refload.export_state(env)
    for o in oracles:
        o.set_state(g[None], p[None], [t])
    rs = np.random.RandomState(seed)
    T = int(2 * R * Cc / ss) + 40
    episodes = 0
    for k in range(T):
        a = rs.randint(0, 4, D)
        ob, r, d = env.step(list(a))
        for o in oracles:
            oo, orr, od = o.step(a[None])
            assert np.array_equal(oo[0], ob.astype(np.float32)), k
            assert float(orr[0]) == float(r) and bool(od[0]) == d, k
            assert np.array_equal(o.pos[0], np.array(env.n_drones_pos, np.float64))
            assert o.meta[0, 0] == env.step_func_count
            assert o.meta[0, 1] == (np.array(env.grid) > 0).sum()          # remaining counter == scan
        if d and k % 3 == 0 and episodes < 4:                               # sometimes reset, sometimes step past done
            env.reset()
            g, p, t = refload.export_state(env)
            for o in oracles:
                o.set_state(g[None], p[None], [t])
            episodes += 1


@pytest.mark.parametrize("seed", [0, 1, 12345, 2**31 + 5])
@pytest.mark.parametrize("R,Cc,D,na", [(10, 10, 3, 5), (5, 5, 3, 5), (32, 32, 4, 5), (64, 64, 16, 5), (1, 7, 2, 3), (256, 3, 5, 9)])
def test_numpy_reset_matches_reference(seed, R, Cc, D, na):
    """oracle_reset_numpy == np.random.seed(seed); DroneEnv(); reset(); reset()  (legacy MT19937 + masked rejection)."""
    kw = dict(row_count=R, col_count=Cc, n_anamolous=na, n_drones=D)
    env = refload.make_ref_env("A2C", seed=seed, **kw)
    for n_resets in (1, 2, 3):
        o = OracleBatch(1, **kw)
        o.reset_numpy(0, seed, n_resets)
        g, p, _ = refload.export_state(env)
        assert np.array_equal(o.grid[0], g) and np.array_equal(o.pos[0], p), (seed, n_resets)
        env.reset()


def _fuzz_once(variant, rm, R, Cc, D, cd, ss, na, U, seed, T):
    kw = dict(row_count=R, col_count=Cc, n_anamolous=na, uncertainity=U, collision_dist=cd, n_drones=D, step_size=ss)
    env = refload.make_ref_env(variant, seed=seed, **kw)
    oracles = [OracleBatch(1, reward_mode=rm, scan_mode=s, **kw) for s in (0, 1)]
    g, p, t = refload.export_state(env)
    for o in oracles:
        o.set_state(g[None], p[None], [t])
    rs = np.random.RandomState(seed ^ 0x5A5A)
    for k in range(T):
        a = rs.randint(0, 4, D)
        ob, r, d = env.step(list(a))
        for o in oracles:
            oo, orr, od = o.step(a[None])
            assert np.array_equal(oo[0], ob.astype(np.float32)), (k, kw)
            assert float(orr[0]) == float(r) and bool(od[0]) == d, (k, kw)
            assert np.array_equal(o.pos[0], np.array(env.n_drones_pos, np.float64)), (k, kw)
            assert o.meta[0, 0] == env.step_func_count


def test_step_parity_fuzz():
    """Seeded random sweep over the constructor space the reference accepts (both reward variants, rectangular grids,
    D > R, every dyadic step_size, collision_dist on both sides of 0.5 and of the lattice diagonals, U != 5): both oracle
    scan modes against the live reference, stepping past `done` and past the punish threshold (noVis_sim.py:141-156)."""
    rs = np.random.RandomState(20261104)
    for case in range(60):
        R, Cc = int(rs.randint(1, 9)), int(rs.randint(1, 9))
        D = int(rs.randint(1, 7))
        cd = float(rs.choice([0.3, 0.4, 0.5, 0.51, 0.75, 1.0, 1.2, 1.42, 1.6, 2.0]))
        ss = float(rs.choice([0.25, 0.5, 1.0, 2.0]))
        na = int(rs.randint(0, min(R * Cc, 6) + 1))
        U = int(rs.choice([2, 5, 9]))
        variant, rm = (("A2C", 0), ("DQL", 1))[case % 2]
        T = int(R * Cc / ss) + 25                                           # crosses t > R*C/ss
        _fuzz_once(variant, rm, R, Cc, D, cd, ss, na, U, 1000 + case, T)
