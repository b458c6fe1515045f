"""GPU: parity at BASELINE.json's FULL sizes against the C oracle (bounded-window scan, itself pinned to the live
reference): rewards / dones every step, the whole observation tensor at a stride, and grid / positions / counters /
statistics of ALL envs at the end.

  cfg2  4096 envs x 4 drones, 32x32, full obs, T = 4096 steps (crosses the punish threshold 1024 and many auto-resets)
  cfg3  65536 envs x 16 drones, 64x64, 5x5 windows, T = 256 steps (the headline shape, LEAN kernel)
  cfg4  16384 envs x 8 drones, 10x10, full obs, T = 1024 (env side of the A2C / deep-Q loops; crosses the punish threshold
        100 and hundreds of auto-resets per env)
  cfg5  per-GPU share of the 8-GPU stress shape: 1024 envs x 64 drones, 256x256 (CTA kernel) — 5a full float obs (268 MB per
        step), 5b 5x5 windows
Actions: np.random.RandomState(1234).randint(0, 4, (T_chunk, N, D)), drawn in chunks (SURVEY §8d synthetic inputs).
"""
import numpy as np
import pytest
import torch

from common import SHAPES
from test_gpu_parity import assert_state_equal, make_pair

pytestmark = pytest.mark.gpu


def lockstep_fullsize(cfg, N, T, obs_every, chunk=32):
    kw = dict(SHAPES[cfg])
    g, o = make_pair(N, auto_reset=True, seed=2026, **kw)
    rs = np.random.RandomState(1234)
    k = 0
    while k < T:
        n = min(chunk, T - k)
        acts = rs.randint(0, 4, (n, N, kw["n_drones"])).astype(np.uint8)
        acts_d = torch.from_numpy(acts).cuda()
        for j in range(n):
            want_obs = (k + j) % obs_every == 0 or k + j == T - 1
            obs, rew, done, _ = g.step(acts_d[j])
            oo, orr, od = o.step(acts[j], want_obs=want_obs)
            assert np.array_equal(rew.cpu().numpy(), orr), f"{cfg} reward, step {k + j}"
            assert np.array_equal(done.cpu().numpy(), od), f"{cfg} done, step {k + j}"
            if want_obs:
                assert np.array_equal(obs.reshape(N, -1).cpu().numpy(), oo), f"{cfg} obs, step {k + j}"
        k += n
    assert_state_equal(g, o, f"{cfg} final")
    assert np.array_equal(g.stats.cpu().numpy(), o.stats)
    assert g.device_error() == 0
    stats = g.stats_dict()
    g.close()
    return stats


def test_cfg2_full_size_full_horizon():
    stats = lockstep_fullsize("cfg2", 4096, 4096, obs_every=256)
    assert stats["env_steps"] == 4096 * 4096 and stats["cells_mapped"] > 0


def test_cfg3_full_size():
    stats = lockstep_fullsize("cfg3", 65536, 256, obs_every=32)
    assert stats["agent_steps"] == 65536 * 16 * 256 and stats["cells_mapped"] > 0


def test_cfg4_full_size():
    stats = lockstep_fullsize("cfg4", 16384, 1024, obs_every=64)
    assert stats["env_steps"] == 16384 * 1024 and stats["episodes_done"] > 16384


def test_cfg5a_per_gpu_size():
    stats = lockstep_fullsize("cfg5a", 1024, 96, obs_every=32)
    assert stats["agent_steps"] == 1024 * 64 * 96 and stats["cells_mapped"] > 0


def test_cfg5b_per_gpu_size():
    stats = lockstep_fullsize("cfg5b", 1024, 512, obs_every=64)
    assert stats["agent_steps"] == 1024 * 64 * 512 and stats["cells_mapped"] > 0
