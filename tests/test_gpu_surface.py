"""GPU: per-stage entry points, rollout, sharding invariance, host-buffer step and the single-env facade."""
import json

import numpy as np
import pytest
import torch

from common import SHAPES, env_kwargs, load_golden
from oracle.oracle import OracleBatch
from test_gpu_parity import assert_state_equal, make_pair

pytestmark = pytest.mark.gpu


def test_stage_entry_points_compose_to_step():
    """K1 move -> K2 scatter -> K3 reward -> K4 obs == the fused step == the oracle."""
    for kw in (dict(SHAPES["cfg2"]), dict(SHAPES["cfg3"]), dict(row_count=8, col_count=5, n_drones=9, collision_dist=1.2)):
        N = 90
        g, o = make_pair(N, seed=13, **kw)
        rs = np.random.RandomState(1)
        for k in range(150):
            a = rs.randint(0, 4, (N, o.D)).astype(np.uint8)
            g.stage_move(torch.from_numpy(a).cuda())
            cover = g.stage_scatter()
            rew, done = g.stage_reward(cover)
            obs = g.stage_obs()
            oo, orr, od = o.step(a)
            assert np.array_equal(obs.reshape(N, -1).cpu().numpy(), oo), k
            assert np.array_equal(rew.cpu().numpy(), orr) and np.array_equal(done.cpu().numpy(), od), k
        assert_state_equal(g, o)
        g.close()


def test_stage_move_clips_like_np_clip():
    g, o = make_pair(64, seed=2, row_count=4, col_count=6, n_drones=5, step_size=0.5)
    for k in range(40):
        a = np.full((64, 5), k % 4, np.uint8)
        g.stage_move(torch.from_numpy(a).cuda())
    p = g.pos.cpu().numpy()
    assert p[..., 0].min() >= 0 and p[..., 0].max() <= 3 and p[..., 1].min() >= 0 and p[..., 1].max() <= 5
    g.close()


@pytest.mark.parametrize("cfg", ["cfg2", "cfg3", "cfg4"])
def test_rollout_equals_steps(cfg):
    kw = dict(SHAPES[cfg])
    N, T = 77, 96
    g1, o = make_pair(N, auto_reset=True, seed=17, **kw)
    g2, _ = make_pair(N, auto_reset=True, seed=17, **kw)
    a = torch.randint(0, 4, (T, N, o.D), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    obs_all = torch.empty((T,) + g1.obs_shape, dtype=torch.float32, device="cuda")
    _, rew, done = g1.rollout(a, obs_out=obs_all)
    for t in range(T):
        ob, r, d, _ = g2.step(a[t])
        assert torch.equal(ob, obs_all[t]), t
        assert torch.equal(r, rew[t]) and torch.equal(d, done[t]), t
        oo, orr, od = o.step(a[t].cpu().numpy())
        assert np.array_equal(oo, obs_all[t].reshape(N, -1).cpu().numpy()), t
    assert_state_equal(g1, o)
    assert torch.equal(g1.stats, g2.stats)
    last = torch.empty(g1.obs_shape, dtype=torch.float32, device="cuda")
    g1.rollout(a, obs_out=last, obs_last_only=True)
    for t in range(T):
        ob, _, _, _ = g2.step(a[t])
    assert torch.equal(last, ob)
    g1.close(); g2.close()


def test_sharding_invariance():
    """Rank g of G owns envs [g*N/G, (g+1)*N/G): identical per-env trajectories for any G (RNG keyed by global env id)."""
    kw = dict(SHAPES["cfg4"])
    N, T = 96, 300
    a = torch.randint(0, 4, (T, N, 8), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(3))
    whole, _ = make_pair(N, auto_reset=True, seed=23, **kw)
    _, rew, done = whole.rollout(a)
    for G in (2, 4, 8):
        n = N // G
        stats = torch.zeros(8, dtype=torch.int64, device="cuda")
        for r in range(G):
            shard, _ = make_pair(n, auto_reset=True, seed=23, env_id_base=r * n, **kw)
            _, rr, dd = shard.rollout(a[:, r * n:(r + 1) * n].contiguous())
            assert torch.equal(rr, rew[:, r * n:(r + 1) * n]) and torch.equal(dd, done[:, r * n:(r + 1) * n])
            gg, gp, gt, _ = shard.get_state()
            wg, wp, wt, _ = whole.get_state(r * n, n)
            assert torch.equal(gg, wg) and torch.equal(gp, wp) and torch.equal(gt, wt)
            stats += shard.stats
            shard.close()
        assert torch.equal(stats, whole.stats), G      # what the NCCL all-reduce of the 64-byte vector produces
    whole.close()


def test_step_host_matches_device_step():
    kw = dict(SHAPES["cfg3"])
    N = 200
    g1, o = make_pair(N, auto_reset=True, seed=4, **kw)
    a_h = torch.empty((N, 16), dtype=torch.uint8).pin_memory()
    obs_h = torch.empty(g1.obs_shape, dtype=torch.float32).pin_memory()
    rew_h = torch.empty(N, dtype=torch.float32).pin_memory()
    done_h = torch.empty(N, dtype=torch.uint8).pin_memory()
    stats_h = torch.zeros(8, dtype=torch.int64).pin_memory()
    rs = np.random.RandomState(0)
    for k in range(50):
        a = rs.randint(0, 4, (N, 16)).astype(np.uint8)
        a_h.copy_(torch.from_numpy(a))
        g1.step_host(a_h, obs_h, rew_h, done_h, stats_h)
        oo, orr, od = o.step(a)
        assert np.array_equal(obs_h.numpy().reshape(N, -1), oo) and np.array_equal(rew_h.numpy(), orr)
        assert np.array_equal(done_h.numpy(), od)
    assert stats_h[0].item() == N and stats_h[1].item() == N * 16
    g1.close()


@pytest.mark.parametrize("name,N", [("cfg3", 4100), ("cfg1", 16389), ("cfg5b", 700), ("cfg2", 1500)])
def test_step_host_large_obs_matches_device_step(name, N):
    """swarm_step_host with observations of 4 MiB and more (ragged env counts, both kernels) against swarm_step on a twin handle.
    (Written for an env-chunked, copy-overlapped form of swarm_step_host that measured no faster and was dropped)"""
    from swarm_b200 import BatchedDroneEnv
    kw = dict(SHAPES[name])
    D = kw["n_drones"]
    host = BatchedDroneEnv(N, auto_reset=True, seed=21, **kw)
    dev = BatchedDroneEnv(N, auto_reset=True, seed=21, **kw)
    assert int(np.prod(host.obs_shape)) * 4 >= 4 << 20
    a_h = torch.empty((N, D), dtype=torch.uint8).pin_memory()
    obs_h = torch.empty(host.obs_shape, dtype=torch.float32).pin_memory()
    rew_h = torch.empty(N, dtype=torch.float32).pin_memory()
    done_h = torch.empty(N, dtype=torch.uint8).pin_memory()
    stats_h = torch.zeros(8, dtype=torch.int64).pin_memory()
    rs = np.random.RandomState(2)
    total = torch.zeros(8, dtype=torch.int64)
    for k in range(12):
        a_h.copy_(torch.from_numpy(rs.randint(0, 4, (N, D)).astype(np.uint8)))
        obs_h.fill_(-7)
        host.step_host(a_h, obs_h, rew_h, done_h, stats_h)
        obs, rew, done, _ = dev.step(a_h.cuda())
        assert torch.equal(obs_h, obs.cpu()), k
        assert torch.equal(rew_h, rew.cpu()) and torch.equal(done_h, done.cpu()), k
        total += stats_h
    assert torch.equal(total, dev.stats.cpu())
    for x, y in zip(host.get_state(), dev.get_state()):
        assert (x is None and y is None) or torch.equal(x, y)
    host.close(); dev.close()


@pytest.mark.parametrize("N,name,pinned,odt", [(1, "cfg1", True, torch.uint8), (7, "cfg4", True, torch.uint8), (16, "cfg3", True, torch.uint8),
                                               (5, "cfg2", False, torch.uint8), (300, "cfg4", True, torch.uint8),
                                               (3, "cfg1", True, torch.float32), (9, "cfg3", True, torch.float32), (2, "cfg5b", True, torch.float32)])
def test_step_host_pos_small_batches_match_oracle(N, name, pinned, odt):
    """swarm_step_host_pos (what the single-env facade calls): small batches hand pinned host buffers to the kernel itself
    (actions read and results written over PCIe), pageable ones and larger batches stage and copy; all equal the oracle, and the
    positions returned equal the resident ones."""
    kw = dict(SHAPES[name])
    D = kw["n_drones"]
    g, o = make_pair(N, auto_reset=True, seed=5, obs_dtype=odt, **kw)
    mk = (lambda t: t.pin_memory()) if pinned else (lambda t: t)
    a_h = mk(torch.empty((N, D), dtype=torch.uint8))
    obs_h = mk(torch.empty(g.obs_shape, dtype=odt))
    rew_h, done_h = mk(torch.empty(N, dtype=torch.float32)), mk(torch.empty(N, dtype=torch.uint8))
    pos_h = mk(torch.empty((N, D, 2), dtype=torch.float32))
    stats_h = mk(torch.zeros(8, dtype=torch.int64))
    rs = np.random.RandomState(1)
    for k in range(120):
        a = rs.randint(0, 4, (N, D)).astype(np.uint8)
        a_h.copy_(torch.from_numpy(a))
        obs_h.fill_(77); pos_h.fill_(-3)
        g.step_host_pos(a_h, obs_h, rew_h, done_h, pos_h, stats_h if k % 2 else None)
        oo, orr, od = o.step(a)
        want = oo.astype(np.int16).astype(np.uint8) if odt == torch.uint8 else oo.astype(np.float32)   # pad -1 -> 255 for uint8
        assert np.array_equal(obs_h.numpy().reshape(N, -1), want), k
        assert np.array_equal(rew_h.numpy(), orr) and np.array_equal(done_h.numpy(), od), k
        assert np.array_equal(pos_h.numpy().astype(np.float64), o.pos), k
        if k % 2:
            assert stats_h[0].item() == N and stats_h[1].item() == N * D
    assert_state_equal(g, o)
    g.close()


def test_greedy_actions_match_oracle_and_finish_episodes():
    g, o = make_pair(33, auto_reset=False, seed=6, row_count=9, col_count=7, n_drones=3, step_size=1.0)
    for k in range(80):
        a = g.greedy_actions()
        assert np.array_equal(a.cpu().numpy(), o.greedy_actions()), k
        g.step(a)
        o.step(a.cpu().numpy())
    assert_state_equal(g, o)
    assert int((g.meta[:, 1] == 0).sum()) == 33          # greedy pursuit maps a 9x7 grid with 3 drones in < 80 steps
    g.close()


# ------------------------------------------------------------------------------------------------ facade
@pytest.mark.parametrize("name", ["cfg1_defaults_a2c", "cfg1_defaults_dql", "trainer_5x5_a2c", "cfg4_10x10_d8"])
def test_facade_reproduces_reference_with_numpy_seed(name):
    """np.random.seed(s); DroneEnv(**kw) -> the facade draws the reference's MT19937 stream in the reference's order, so
    construction, reset-after-done and every step equal the golden trace of the live reference (3-tuple API)."""
    from swarm_b200 import DQLDroneEnv, DroneEnv
    gd = load_golden(name)
    cfg = gd["config"]
    kw, rm = env_kwargs(cfg)
    np.random.seed(cfg["seed"])
    env = (DroneEnv if rm == 0 else DQLDroneEnv)(**kw)
    assert np.array_equal(np.array(env.grid, np.int8).T, gd["grid0"])
    assert np.array_equal(np.array(env.n_drones_pos, np.float64), gd["pos0"])
    assert env.action_size == 4 and env.state_size == kw["row_count"] * kw["col_count"]
    T = gd["actions"].shape[0]
    for k in range(T):
        acts = [np.array([a]) for a in gd["actions"][k]] if k % 2 else [int(a) for a in gd["actions"][k]]   # A2C/train.py:73 vs :260
        obs, r, d = env.step(acts)
        assert obs.dtype == np.int64 and obs.shape == (env.state_size,)
        assert np.array_equal(obs.astype(np.int8), gd["obs"][k]), k
        assert isinstance(r, float) and r == gd["reward"][k] and isinstance(d, bool) and d == bool(gd["done"][k]), k
        assert np.array_equal(np.array(env.n_drones_pos, np.float64), gd["pos"][k])
        assert env.step_func_count == gd["t"][k]
        if d and cfg["reset_on_done"]:
            ob = env.reset()
            i = list(gd["reset_steps"]).index(k)
            assert np.array_equal(ob.astype(np.int8).reshape(kw["col_count"], kw["row_count"]), gd["reset_grid"][i])
            assert np.array_equal(np.array(env.n_drones_pos, np.float64), gd["reset_pos"][i])
    flags = env.uncertain_points
    assert all(flags[k] == (1 if env.grid[k[0]][k[1]] != 0 else 0) for k in flags)
    env.close()


def test_facade_asserts_like_reference():
    from swarm_b200 import DroneEnv
    env = DroneEnv()
    with pytest.raises(AssertionError, match="provide actions for each drone"):
        env.step([0, 1])
    with pytest.raises(AssertionError, match="action should be in range"):
        env.step([0, 1, 4])
    env.close()


def test_launcher_installs_facade_module():
    import sys
    from swarm_b200 import launcher
    saved = sys.modules.pop("noVis_sim", None)
    try:
        for variant, mode in (("A2C", "A2C"), ("DQL", "DQL")):
            mod = launcher.install(variant)
            from noVis_sim import DroneEnv
            assert DroneEnv is mod.DroneEnv and DroneEnv.reward_mode == mode
            del sys.modules["noVis_sim"]
    finally:
        if saved is not None:
            sys.modules["noVis_sim"] = saved


def test_sharding_invariance_on_one_gpu():
    """Two shards keyed by global env id (env_id_base) reproduce the unsharded batch env for env."""
    import numpy as np
    from swarm_b200 import BatchedDroneEnv, shard_range
    kw = dict(row_count=16, col_count=16, n_drones=8, step_size=1.0, auto_reset=True, seed=123, obs_mode="window", window=5)
    total, T = 301, 150
    full = BatchedDroneEnv(total, **kw)
    shards = []
    for r in range(3):
        first, count = shard_range(total, r, 3)
        shards.append((first, count, BatchedDroneEnv(count, env_id_base=first, **kw)))
    acts = torch.randint(0, 4, (T, total, 8), dtype=torch.uint8, device="cuda")
    for t in range(T):
        of, rf, df, _ = full.step(acts[t])
        for first, count, e in shards:
            o, r, d, _ = e.step(acts[t, first:first + count].contiguous())
            assert torch.equal(o, of[first:first + count]) and torch.equal(r, rf[first:first + count])
            assert torch.equal(d, df[first:first + count])
    tot = sum(e.stats for _, _, e in shards)
    assert torch.equal(tot, full.stats)
    for first, count, e in shards:
        assert torch.equal(e.pos, full.pos[first:first + count]) and torch.equal(e.meta, full.meta[first:first + count])


def test_stats_allreduce_through_library_communicator():
    """swarm_comm_* + swarm_stats_allreduce on a world of one rank (the N > 1 case runs under torchrun in bench.py --gpus N):
    the NCCL call path works, is stream-ordered and leaves a single rank's statistics unchanged."""
    from swarm_b200 import BatchedDroneEnv, StatsComm
    env = BatchedDroneEnv(64, row_count=8, col_count=8, n_drones=4, step_size=1.0, seed=5)
    a = torch.randint(0, 4, (64, 4), dtype=torch.uint8, device="cuda")
    for _ in range(5):
        env.step(a)
    want = env.stats.clone()
    comm = StatsComm()
    assert comm.world == 1
    got = comm.allreduce(env.stats.clone())
    torch.cuda.synchronize()
    assert torch.equal(got, want) and int(want[0]) == 5 * 64
    comm.close()
    env.close()


@pytest.mark.parametrize("cfg", ["cfg2", "cfg3", "cfg4"])
def test_lean_and_generic_kernels_agree(cfg):
    """The LEAN instantiation of the staged kernel (full CTA groups, D == tile, one step per launch) and the generic FAST one
    (forced with SWARM_NO_LEAN=1, read at every launch) produce identical obs / rewards / dones / state / statistics."""
    import os
    from swarm_b200 import BatchedDroneEnv
    kw = dict(SHAPES[cfg])
    om = kw.pop("obs_mode", 0)
    kw["obs_mode"] = "window" if om == 1 else "full"
    N = 2048
    a = BatchedDroneEnv(N, auto_reset=True, seed=9, **kw)
    b = BatchedDroneEnv(N, auto_reset=True, seed=9, **kw)
    rs = np.random.RandomState(2)
    try:
        for k in range(40):
            act = torch.from_numpy(rs.randint(0, 4, (N, kw["n_drones"])).astype(np.uint8)).cuda()
            os.environ.pop("SWARM_NO_LEAN", None)
            oa, ra, da, _ = a.step(act)
            os.environ["SWARM_NO_LEAN"] = "1"
            ob, rb, db, _ = b.step(act)
            assert torch.equal(oa, ob) and torch.equal(ra, rb) and torch.equal(da, db), f"step {k}"
    finally:
        os.environ.pop("SWARM_NO_LEAN", None)
    ga, gb = a.get_state(), b.get_state()
    assert all(torch.equal(x, y) for x, y in zip(ga[:3], gb[:3])) and torch.equal(a.stats, b.stats)
    a.close(); b.close()


def test_two_live_envs_of_one_instantiation_keep_their_shared_memory_limit():
    """the dynamic shared-memory limit belongs to the kernel function, not to a handle.  Env A (large slice) steps,
    env B (same instantiation, smaller slice) takes its first step, then A steps again — A's launch must still be accepted."""
    from test_gpu_parity import make_pair, run_lockstep
    big = dict(row_count=64, col_count=64, n_drones=16, step_size=1.0, obs_mode=1, window=5)
    small = dict(row_count=16, col_count=16, n_drones=16, step_size=1.0, obs_mode=1, window=5)
    ga, oa = make_pair(64, auto_reset=True, seed=5, **big)
    run_lockstep(ga, oa, 4, seed=1)
    gb, ob = make_pair(64, auto_reset=True, seed=6, **small)
    run_lockstep(gb, ob, 4, seed=2)
    run_lockstep(ga, oa, 4, seed=3)          # A again, after B's first launch
    run_lockstep(gb, ob, 4, seed=4)
    for a, b in ((dict(row_count=32, col_count=32, n_drones=4, step_size=1.0), dict(row_count=8, col_count=8, n_drones=4, step_size=1.0)),
                 (dict(row_count=256, col_count=256, n_drones=64, step_size=1.0), dict(row_count=40, col_count=40, n_drones=64, step_size=1.0))):
        g1, o1 = make_pair(48, auto_reset=True, seed=7, **a)
        run_lockstep(g1, o1, 3, seed=1)
        g2, o2 = make_pair(48, auto_reset=True, seed=8, **b)
        run_lockstep(g2, o2, 3, seed=2)
        run_lockstep(g1, o1, 3, seed=3)
        g1.close(); g2.close()
    ga.close(); gb.close()


@pytest.mark.parametrize("kw", [dict(row_count=2, col_count=2, n_drones=2, n_anamolous=1, step_size=1.0),
                                dict(row_count=1, col_count=4, n_drones=3, n_anamolous=1, step_size=1.0),
                                dict(row_count=4, col_count=1, n_drones=1, n_anamolous=0, step_size=0.5),
                                dict(row_count=2, col_count=3, n_drones=5, n_anamolous=2, step_size=1.0)])
def test_tiny_grids_full_obs(kw):
    """R*C == 4 made the (env, quad) magic divisor overflow, so every env after the first of a warp group read
    the wrong words.  2x2 / 1x4 / 4x1 (and a 6-cell grid) against the oracle, many envs per warp."""
    from test_gpu_parity import make_pair, run_lockstep
    g, o = make_pair(77, auto_reset=True, seed=13, **kw)
    run_lockstep(g, o, 60, seed=5)
    g.close()


def test_stats_timeout_is_reported_not_summed():
    """VERDICT r1 / ADVICE r1: a mailbox slot that never becomes consistent must surface as an error, not as a torn sum."""
    from swarm_b200 import BatchedDroneEnv, SwarmError
    env = BatchedDroneEnv(64, row_count=8, col_count=8, n_drones=4, step_size=1.0, auto_reset=True, seed=1).enable_peer_stats()
    env.step(torch.zeros((64, 4), dtype=torch.uint8, device=env.device))
    env.push_stats()
    assert env.global_stats_host(min_pushes=1)[0] == 64
    with pytest.raises(SwarmError, match="consistent"):
        env.global_stats_host(min_pushes=10 ** 9)        # nobody will ever push that often: bounded wait, then an error
    assert env.device_error() & 4
    g = env.global_stats(min_pushes=10 ** 9)
    torch.cuda.synchronize()
    assert g.tolist() == [-1] * 8
    assert env.device_error() & 4
    env.close()
