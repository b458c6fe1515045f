"""GPU: the CUDA path (through the C ABI) against the CPU oracle and the golden traces of the live reference.
Bit-exact for grid / obs / done / counters; rewards and positions exact (integers / dyadic lattice) — the 1e-6 relative
tolerance north_star allows is written below only as a backstop."""
import numpy as np
import pytest
import torch

from common import SHAPES, env_kwargs, golden_names, load_golden
from oracle.oracle import OracleBatch

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def make_pair(N, auto_reset=False, reward_mode=0, obs_mode=0, window=5, seed=7, env_id_base=0, dwell=(0, 0),
              obs_dtype=None, **kw):
    from swarm_b200 import BatchedDroneEnv
    base = dict(row_count=10, col_count=10, n_anamolous=5, uncertainity=5, collision_dist=0.5, n_drones=3, step_size=0.5)
    base.update(kw)
    g = BatchedDroneEnv(N, reward_mode=reward_mode, obs_mode=obs_mode, window=window, auto_reset=auto_reset, seed=seed,
                        env_id_base=env_id_base, dwell_normal=dwell[0], dwell_anom=dwell[1],
                        obs_dtype=obs_dtype or torch.float32, **base)
    o = OracleBatch(N, reward_mode=reward_mode, obs_mode=obs_mode, window=window, auto_reset=auto_reset, seed=seed,
                    env_id_base=env_id_base, dwell_normal=dwell[0], dwell_anom=dwell[1], scan_mode=1, nthreads=0, **base)
    return g, o


def assert_state_equal(g, o, ctx=""):
    gg, gp, gt, gw = g.get_state()
    assert np.array_equal(gg.cpu().numpy(), o.grid), f"grid mismatch {ctx}"
    np.testing.assert_allclose(gp.cpu().numpy().astype(np.float64), o.pos, rtol=RTOL, atol=0, err_msg=f"pos {ctx}")
    assert np.array_equal(g.meta.cpu().numpy(), o.meta), f"meta mismatch {ctx}"
    if o.wait is not None:
        assert np.array_equal(gw.cpu().numpy(), o.wait), f"wait mismatch {ctx}"


def run_lockstep(g, o, T, seed=0, check_every=1, obs_cast=np.float32):
    rs = np.random.RandomState(seed)
    N, D = o.N, o.D
    for k in range(T):
        a = rs.randint(0, 4, (N, D)).astype(np.uint8)
        obs, rew, done, _ = g.step(torch.from_numpy(a).cuda())
        oo, orr, od = o.step(a)
        if k % check_every == 0 or k == T - 1:
            assert np.array_equal(obs.reshape(N, -1).cpu().numpy().astype(obs_cast), oo.astype(obs_cast)), f"obs step {k}"
            np.testing.assert_allclose(rew.cpu().numpy(), orr, rtol=RTOL, atol=0, err_msg=f"reward step {k}")
            assert np.array_equal(done.cpu().numpy(), od), f"done step {k}"
    assert_state_equal(g, o, "final")
    assert np.array_equal(g.stats.cpu().numpy(), o.stats), (g.stats.cpu().numpy(), o.stats)
    assert g.device_error() == 0


# ------------------------------------------------------------------------------------------------ initial state
def test_create_matches_oracle_init():
    for kw in (dict(), dict(row_count=32, col_count=32, n_drones=4), dict(row_count=7, col_count=13, n_drones=40, n_anamolous=60),
               dict(row_count=64, col_count=64, n_drones=16)):
        g, o = make_pair(257, seed=99, env_id_base=1000, **kw)
        assert_state_equal(g, o, str(kw))
        assert np.array_equal(g.stage_obs().cpu().numpy(), o.obs())
        g.close()


def test_reset_mask_and_episode_keys():
    g, o = make_pair(130, seed=5)
    mask = (np.arange(130) % 3 == 0)
    obs = g.reset(torch.from_numpy(mask).cuda())
    oo = o.reset(mask)
    assert np.array_equal(obs.cpu().numpy(), oo)
    assert_state_equal(g, o)
    obs = g.reset()
    oo = o.reset()
    assert np.array_equal(obs.cpu().numpy(), oo)
    assert_state_equal(g, o)


# ------------------------------------------------------------------------------------------------ golden traces
@pytest.mark.parametrize("name", golden_names())
def test_cuda_matches_reference_golden(name):
    """N copies of the golden env (injected state), stepped with the golden actions: obs/reward/done/pos per step."""
    from swarm_b200 import BatchedDroneEnv
    gd = load_golden(name)
    kw, rm = env_kwargs(gd["config"])
    N = 5
    env = BatchedDroneEnv(N, reward_mode=rm, auto_reset=False, **kw)
    rep = lambda a: np.repeat(np.asarray(a)[None], N, 0)
    env.set_state(rep(gd["grid0"]), rep(gd["pos0"]), np.full(N, int(gd["t0"]), np.int32))
    resets = {int(s): i for i, s in enumerate(gd["reset_steps"])}
    T = gd["actions"].shape[0]
    acts = torch.from_numpy(np.repeat(gd["actions"][:, None, :], N, 1)).cuda()
    for k in range(T):
        obs, rew, done, _ = env.step(acts[k])
        o = obs.cpu().numpy()
        assert np.array_equal(o.astype(np.int8), rep(gd["obs"][k])), (name, k)
        assert np.all(rew.cpu().numpy().astype(np.float64) == gd["reward"][k]), (name, k)
        assert np.all(done.cpu().numpy() == gd["done"][k]), (name, k)
        assert np.array_equal(env.pos.cpu().numpy().astype(np.float64), rep(gd["pos"][k])), (name, k)
        assert np.all(env.meta[:, 0].cpu().numpy() == gd["t"][k])
        if k in resets:
            i = resets[k]
            env.set_state(rep(gd["reset_grid"][i]), rep(gd["reset_pos"][i]), np.zeros(N, np.int32))
    env.close()


# ------------------------------------------------------------------------------------------------ lockstep vs oracle
@pytest.mark.parametrize("cfg", ["cfg1", "cfg2", "cfg3", "cfg4"])
@pytest.mark.parametrize("auto_reset", [False, True])
def test_lockstep_baseline_shapes(cfg, auto_reset):
    kw = dict(SHAPES[cfg])
    N = {"cfg1": 300, "cfg2": 130, "cfg3": 67, "cfg4": 300}[cfg]
    T = {"cfg1": 700, "cfg2": 400, "cfg3": 200, "cfg4": 500}[cfg]
    g, o = make_pair(N, auto_reset=auto_reset, seed=11, **kw)
    run_lockstep(g, o, T, seed=3, check_every=7)
    g.close()


def test_lockstep_cfg5_shapes():
    for cfg, N, T in (("cfg5a", 9, 40), ("cfg5b", 9, 60)):
        g, o = make_pair(N, auto_reset=True, seed=2, **SHAPES[cfg])
        run_lockstep(g, o, T, seed=4, check_every=5)
        g.close()


@pytest.mark.parametrize("kw", [
    dict(row_count=5, col_count=5, n_drones=3, step_size=1.0, reward_mode=1),
    dict(row_count=6, col_count=9, n_drones=4, step_size=1.0, collision_dist=0.4),
    dict(row_count=8, col_count=5, n_drones=9, step_size=0.5, collision_dist=1.2),
    dict(row_count=7, col_count=7, n_drones=2, step_size=0.25, collision_dist=1.6, reward_mode=1),
    dict(row_count=12, col_count=12, n_drones=40, step_size=1.0, n_anamolous=7),
    dict(row_count=4, col_count=4, n_drones=1, step_size=1.0, n_anamolous=1),
    dict(row_count=3, col_count=17, n_drones=33, step_size=1.0, n_anamolous=0),
    dict(row_count=16, col_count=16, n_drones=32, step_size=1.0, obs_mode=1, window=3),
    dict(row_count=9, col_count=11, n_drones=5, step_size=0.5, obs_mode=1, window=7),
    dict(row_count=20, col_count=20, n_drones=70, step_size=1.0, obs_mode=1, window=5, collision_dist=1.2),
])
@pytest.mark.parametrize("auto_reset", [False, True])
def test_lockstep_edge_shapes(kw, auto_reset):
    g, o = make_pair(101, auto_reset=auto_reset, seed=21, **kw)
    run_lockstep(g, o, 250, seed=8, check_every=3)
    g.close()


@pytest.mark.parametrize("kw", [
    dict(row_count=6, col_count=6, n_drones=6, step_size=1.0, dwell=(3, 7)),
    dict(row_count=5, col_count=5, n_drones=12, step_size=1.0, dwell=(10, 35)),
    dict(row_count=8, col_count=8, n_drones=40, step_size=0.5, dwell=(2, 5)),
    dict(row_count=6, col_count=6, n_drones=4, step_size=1.0, dwell=(2, 3), collision_dist=1.2),
    # the CTA kernel's all-thread path with dwell: groups of drones on one cell that span several warps, and the W = 5 window path
    dict(row_count=3, col_count=3, n_drones=64, step_size=1.0, dwell=(10, 35)),
    dict(row_count=4, col_count=5, n_drones=100, step_size=1.0, dwell=(7, 20), n_anamolous=3),
    dict(row_count=40, col_count=40, n_drones=48, step_size=1.0, obs_mode=1, window=5, dwell=(2, 5)),
    dict(row_count=12, col_count=9, n_drones=33, step_size=1.0, obs_mode=1, window=3, dwell=(1, 2), n_anamolous=12),
])
def test_lockstep_dwell(kw):
    """Dwell counters (swarmsim-2d/goals.py semantics; ours, parity unpinned vs the reference)."""
    g, o = make_pair(64, auto_reset=True, seed=31, **kw)
    run_lockstep(g, o, 400, seed=9, check_every=3)
    g.close()


def test_uint8_obs():
    for kw in (dict(SHAPES["cfg2"]), dict(SHAPES["cfg3"])):
        g, o = make_pair(50, auto_reset=True, seed=3, obs_dtype=torch.uint8, **kw)
        rs = np.random.RandomState(0)
        for k in range(60):
            a = rs.randint(0, 4, (50, o.D)).astype(np.uint8)
            obs, _, _, _ = g.step(torch.from_numpy(a).cuda())
            oo, _, _ = o.step(a)
            assert np.array_equal(obs.reshape(50, -1).cpu().numpy(), oo.astype(np.int16).astype(np.uint8)), k   # -1 -> 255
        g.close()


@pytest.mark.parametrize("kw,N", [
    (dict(row_count=64, col_count=64, n_drones=16, step_size=1.0), 130),                     # cfg3: LEAN instantiation
    (dict(row_count=64, col_count=64, n_drones=16, step_size=1.0), 131),                     # ragged: FAST, partial last group
    (dict(row_count=10, col_count=10, n_drones=3, step_size=0.5), 57),                        # fractional positions, pads everywhere
    (dict(row_count=7, col_count=33, n_drones=8, step_size=1.0, n_anamolous=8), 64),          # rows not word-aligned
    (dict(row_count=32, col_count=32, n_drones=32, step_size=1.0, collision_dist=1.2), 40),   # generic instantiation
    (dict(row_count=12, col_count=12, n_drones=5, step_size=1.0, dwell=(2, 3)), 33),
])
def test_packed2_windows_match_oracle(kw, N):
    """obs_dtype="packed2" (SWARM_DTYPE_P2): one 64-bit word per drone = the two bit planes of its 5x5 window.  Expanded with
    unpack_windows it must equal the oracle's float windows (and the uint8 form) at every step; rewards / dones / state too."""
    from swarm_b200 import unpack_windows
    kw = dict(kw)
    g, o = make_pair(N, auto_reset=True, seed=23, obs_mode=1, window=5, obs_dtype=torch.int64, **kw)
    assert g.obs_shape == (N, o.D)
    rs = np.random.RandomState(6)
    for k in range(150):
        a = rs.randint(0, 4, (N, o.D)).astype(np.uint8)
        obs, rew, done, _ = g.step(torch.from_numpy(a).cuda())
        oo, orr, od = o.step(a)
        assert obs.dtype == torch.int64 and tuple(obs.shape) == (N, o.D)
        vals = unpack_windows(obs, o.cfg.uncertainity if hasattr(o, "cfg") else 5).reshape(N, -1).cpu().numpy()
        assert np.array_equal(vals, oo.astype(np.float32)), k
        assert np.array_equal(rew.cpu().numpy(), orr.astype(np.float32)) and np.array_equal(done.cpu().numpy(), od), k
    u8 = unpack_windows(obs, 5, torch.uint8).reshape(N, -1).cpu().numpy()
    assert np.array_equal(u8, oo.astype(np.int16).astype(np.uint8))
    assert_state_equal(g, o)
    assert np.array_equal(g.stats.cpu().numpy(), o.stats)
    # rollout and the stage entry point emit the same format
    T = 4
    a = rs.randint(0, 4, (T, N, o.D)).astype(np.uint8)
    out = torch.empty((T, N, o.D), dtype=torch.int64, device="cuda")
    g.rollout(torch.from_numpy(a).cuda(), obs_out=out)
    for t in range(T):
        oo, _, _ = o.step(a[t])
        assert np.array_equal(unpack_windows(out[t], 5).reshape(N, -1).cpu().numpy(), oo.astype(np.float32)), t
    assert np.array_equal(unpack_windows(g.stage_obs(), 5).reshape(N, -1).cpu().numpy(), oo.astype(np.float32))
    # host-buffer step
    obs_h = torch.empty((N, o.D), dtype=torch.int64).pin_memory()
    a_h = torch.from_numpy(rs.randint(0, 4, (N, o.D)).astype(np.uint8)).pin_memory()
    rew_h, done_h = torch.empty(N, dtype=torch.float32).pin_memory(), torch.empty(N, dtype=torch.uint8).pin_memory()
    g.step_host(a_h, obs_h, rew_h, done_h)
    oo, orr, od = o.step(a_h.numpy())
    assert np.array_equal(unpack_windows(obs_h, 5).reshape(N, -1).numpy(), oo.astype(np.float32)) and np.array_equal(rew_h.numpy(), orr.astype(np.float32))
    g.close()


def test_packed2_needs_the_staged_kernel():
    from swarm_b200 import BatchedDroneEnv, SwarmError
    with pytest.raises(SwarmError):          # D > 32: one-CTA-per-env kernel, no packed emitter
        BatchedDroneEnv(8, row_count=64, col_count=64, n_drones=40, step_size=1.0, obs_mode="window", window=5, obs_dtype="packed2")
    with pytest.raises(SwarmError):          # more than 8 anomalies: PLANE mode
        BatchedDroneEnv(8, row_count=32, col_count=32, n_drones=4, n_anamolous=12, step_size=1.0, obs_mode="window", window=5, obs_dtype="packed2")
    g = BatchedDroneEnv(8, row_count=32, col_count=32, n_drones=4, step_size=1.0, obs_mode="window", window=5, obs_dtype="packed2")
    assert g.obs_dtype == torch.int64 and g.obs_shape == (8, 4)
    g.close()


def test_action_dtypes_and_bad_action():
    g, o = make_pair(40, seed=1, **SHAPES["cfg4"])
    rs = np.random.RandomState(5)
    for dt in (torch.uint8, torch.int32, torch.int64):
        a = rs.randint(0, 4, (40, 8))
        g.step(torch.from_numpy(a).cuda().to(dt))
        o.step(a.astype(np.uint8))
    assert_state_equal(g, o)
    assert g.device_error() == 0
    a = torch.zeros((40, 8), dtype=torch.int64, device="cuda")
    a[3, 2] = 4
    g.step(a)
    assert g.device_error() == 1          # sticky word, cleared by the read
    assert g.device_error() == 0
    with pytest.raises(AssertionError):
        g.step(torch.zeros((40, 7), dtype=torch.uint8, device="cuda"))
    g.close()


@pytest.mark.parametrize("name", ["cfg2", "cfg3", "cfg4"])
def test_int_action_dtypes_take_the_fast_kernel_and_match(name):
    """int32 / int64 actions (what torch's Categorical.sample() returns) run the FAST instantiation with direct action loads:
    obs / reward / done / state equal the oracle's, step for step, negative and huge int64 codes included (error word)."""
    kw = dict(SHAPES[name])
    N, D = 333, kw["n_drones"]
    g, o = make_pair(N, auto_reset=True, seed=17, **kw)
    rs = np.random.RandomState(4)
    for k in range(90):
        a = rs.randint(0, 4, (N, D))
        dt = (torch.int64, torch.int32, torch.uint8)[k % 3]
        obs, rew, done, _ = g.step(torch.from_numpy(a).cuda().to(dt))
        oo, orr, od = o.step(a.astype(np.uint8))
        assert np.array_equal(obs.reshape(N, -1).cpu().numpy(), oo), (k, dt)
        assert np.array_equal(rew.cpu().numpy(), orr) and np.array_equal(done.cpu().numpy(), od), (k, dt)
    assert_state_equal(g, o)
    assert g.device_error() == 0
    before = g.pos.clone()
    bad = torch.full((N, D), -1, dtype=torch.int64, device="cuda")
    bad[0, 0] = 2 ** 40 + 1                    # the low byte alone would be a valid code
    g.step(bad)
    assert g.device_error() == 1
    assert torch.equal(g.pos, before)          # an invalid code moves nothing
    g.close()
