"""GPU: the peer-memory form of the statistics exchange (swarm_stats_mailbox_handle / peers_open / global): every step launch
publishes the rank's running totals (as of the previous launch) into every rank's mailbox; swarm_stats_push is the exact flush.  World of one on a single GPU; two processes over
NVLink when the box has two GPUs (the round-end box has one: the 2-rank case then runs only in bench.py --gpus N)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _step_some(env, n_steps, seed):
    rs = np.random.RandomState(seed)
    for _ in range(n_steps):
        env.step(torch.from_numpy(rs.randint(0, 4, (env.num_envs, env.n_drones)).astype(np.uint8)).to(env.device))


@pytest.mark.parametrize("kw", [dict(row_count=64, col_count=64, n_drones=16, step_size=1.0, obs_mode="window"),      # staged kernel
                                dict(row_count=10, col_count=10, n_drones=3, step_size=0.5),
                                dict(row_count=256, col_count=256, n_drones=64, step_size=1.0, obs_mode="window")])   # CTA kernel
def test_world_of_one_mailbox_tracks_running_totals(kw):
    from swarm_b200 import BatchedDroneEnv
    env = BatchedDroneEnv(96, auto_reset=True, seed=4, **kw).enable_peer_stats()
    for k in range(1, 6):
        before = env.stats.clone()
        _step_some(env, 3, seed=k)
        lag = env.global_stats(min_pushes=4 * k - 1).clone()    # the in-step feed: totals as the launch before the last left them
        env.push_stats()                                         # explicit, exact
        g = env.global_stats(min_pushes=4 * k)
        torch.cuda.synchronize()
        assert torch.equal(g, env.stats), (k, g.tolist(), env.stats.tolist())
        assert int(before[0]) <= int(lag[0]) <= int(env.stats[0]) and int(lag[0]) == int(env.stats[0]) - 96
    assert int(env.stats[0]) == 15 * 96
    env.disable_peer_stats()
    _step_some(env, 2, seed=9)          # pushes are off again; stepping still works
    assert int(env.stats[0]) == 17 * 96
    env.close()


def _worker(rank, world, port, steps, out_q):
    import torch.distributed as dist
    from swarm_b200 import BatchedDroneEnv
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    env = BatchedDroneEnv(512, row_count=16, col_count=16, n_drones=4, step_size=1.0, auto_reset=True, seed=3, env_id_base=rank * 512)
    env.enable_peer_stats()
    _step_some(env, steps, seed=100 + rank)
    env.push_stats()
    torch.cuda.synchronize()
    dist.barrier()
    g = env.global_stats(min_pushes=steps + 1)
    torch.cuda.synchronize()
    tot = env.stats.cpu().clone()
    dist.all_reduce(tot)
    out_q.put((rank, g.cpu().tolist(), tot.tolist()))
    dist.barrier()
    env.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs of one node")
def test_two_ranks_push_over_peer_memory():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29611, 12, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, got, want in res:
        assert got == want, (rank, got, want)
