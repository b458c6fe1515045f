"""GPU: swarm_step_host_static (the allocation-free host step: staging in the library's static device array) against the
existing swarm_step_host on a twin handle, bit for bit, on consecutive calls.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest
import torch

from common import SHAPES
from swarm_b200 import BatchedDroneEnv, _lib

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 4 * 1024 * 1024       # SWARM_STATIC_SCRATCH_BYTES, include/swarm_b200.h


def host_buffers(g, N, D):
    pin = lambda t: t.pin_memory()
    return dict(a=pin(torch.empty((N, D), dtype=torch.uint8)), obs=pin(torch.empty(g.obs_shape, dtype=g.obs_dtype)),
                rew=pin(torch.empty(N, dtype=torch.float32)), done=pin(torch.empty(N, dtype=torch.uint8)),
                stats=pin(torch.zeros(8, dtype=torch.int64)))


def vp(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


@pytest.mark.parametrize("name,N", [("cfg1", 512), ("cfg3", 200), ("cfg4", 1000)])
def test_static_scratch_step_equals_step_host(name, N):
    kw = dict(SHAPES[name])
    D = kw["n_drones"]
    L = _lib.lib()
    ref = BatchedDroneEnv(N, auto_reset=True, seed=11, **kw)          # existing entry point
    new = BatchedDroneEnv(N, auto_reset=True, seed=11, **kw)          # new entry point, same envs
    hr, hn = host_buffers(ref, N, D), host_buffers(new, N, D)
    rs = np.random.RandomState(3)
    free_before = None
    for k in range(8):                                                # consecutive calls share (and reuse) the static array
        a = torch.from_numpy(rs.randint(0, 4, (N, D)).astype(np.uint8))
        hr["a"].copy_(a); hn["a"].copy_(a)
        ref.step_host(hr["a"], hr["obs"], hr["rew"], hr["done"], hr["stats"])
        if free_before is None:
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        hn["obs"].fill_(7); hn["rew"].fill_(-5); hn["done"].fill_(9); hn["stats"].fill_(-1)
        rc = L.swarm_step_host_static(new._h, vp(hn["a"]), vp(hn["obs"]), vp(hn["rew"]), vp(hn["done"]), vp(hn["stats"]), None)
        assert rc == 0, L.swarm_last_cuda_error()
        assert torch.equal(hn["obs"], hr["obs"]), k                   # integer-valued obs / rewards: bit-exact
        assert torch.equal(hn["rew"], hr["rew"]) and torch.equal(hn["done"], hr["done"]), k
        assert torch.equal(hn["stats"], hr["stats"]), k               # per-call statistics: the slot was cleared
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before                # the calls allocated no device memory
    for x, y in zip(ref.get_state(), new.get_state()):
        assert (x is None and y is None) or torch.equal(x, y)
    # obs_host == NULL (observation not wanted): reward / done / statistics still match
    a = torch.from_numpy(rs.randint(0, 4, (N, D)).astype(np.uint8))
    hr["a"].copy_(a); hn["a"].copy_(a)
    ref.step_host(hr["a"], None, hr["rew"], hr["done"], hr["stats"])
    new.step_host_static(hn["a"], None, hn["rew"], hn["done"], hn["stats"])
    assert torch.equal(hn["rew"], hr["rew"]) and torch.equal(hn["done"], hr["done"]) and torch.equal(hn["stats"], hr["stats"])
    ref.close(); new.close()


def test_static_scratch_two_handles_share_the_array():
    """Two handles of different shapes alternate on the one array (serialised by the call's own stream sync)."""
    ka, kb = dict(SHAPES["cfg1"]), dict(SHAPES["cfg2"])
    Na, Nb = 300, 77
    pairs = []
    for kw, N in ((ka, Na), (kb, Nb)):
        ref = BatchedDroneEnv(N, auto_reset=True, seed=5, **kw)
        new = BatchedDroneEnv(N, auto_reset=True, seed=5, **kw)
        pairs.append((ref, new, host_buffers(ref, N, kw["n_drones"]), host_buffers(new, N, kw["n_drones"]), N, kw["n_drones"]))
    rs = np.random.RandomState(8)
    for k in range(6):
        for ref, new, hr, hn, N, D in pairs:
            a = torch.from_numpy(rs.randint(0, 4, (N, D)).astype(np.uint8))
            hr["a"].copy_(a); hn["a"].copy_(a)
            ref.step_host(hr["a"], hr["obs"], hr["rew"], hr["done"], hr["stats"])
            new.step_host_static(hn["a"], hn["obs"], hn["rew"], hn["done"], hn["stats"])
            for key in ("obs", "rew", "done", "stats"):
                assert torch.equal(hn[key], hr[key]), (k, key)
    for ref, new, *_ in pairs:
        ref.close(); new.close()


def test_static_scratch_rejects_oversize():
    kw = dict(SHAPES["cfg3"])
    N = 4096                                                          # obs float32 [4096][16][25] = 6.5 MB > 4 MiB
    g = BatchedDroneEnv(N, auto_reset=True, seed=1, **kw)
    assert int(np.prod(g.obs_shape)) * 4 > STATIC_SCRATCH_BYTES
    h = host_buffers(g, N, 16)
    h["a"].zero_()
    L = _lib.lib()
    before = g.get_state()
    rc = L.swarm_step_host_static(g._h, vp(h["a"]), vp(h["obs"]), vp(h["rew"]), vp(h["done"]), vp(h["stats"]), None)
    assert rc == -1
    msg = L.swarm_last_cuda_error().decode()
    assert "static scratch" in msg and str(STATIC_SCRATCH_BYTES) in msg, msg
    with pytest.raises(_lib.SwarmError, match="static scratch"):
        g.step_host_static(h["a"], h["obs"], h["rew"], h["done"], h["stats"])
    for x, y in zip(before, g.get_state()):                            # a rejected call does not touch the envs
        assert (x is None and y is None) or torch.equal(x, y)
    # without the observation the same handle fits (actions 64 KB + reward 16 KB + done 4 KB)
    g.step_host_static(h["a"], None, h["rew"], h["done"], h["stats"])
    assert h["stats"][0].item() == N and h["stats"][1].item() == N * 16
    assert L.swarm_step_host_static(None, vp(h["a"]), None, vp(h["rew"]), vp(h["done"]), None, None) == -1
    g.close()
