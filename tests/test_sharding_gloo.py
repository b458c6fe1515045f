"""N > 1 host logic on CPU: world_size-2 gloo processes, each stepping its shard (the CPU oracle stands in for the device
env), must reproduce the unsharded run env for env, and the statistics all-reduce must equal the unsharded totals."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle.oracle import OracleBatch
from swarm_b200.sharding import reduce_stats, shard_range

KW = dict(row_count=8, col_count=8, n_drones=4, step_size=1.0, auto_reset=True, seed=77, scan_mode=1)
TOTAL, T = 37, 120


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    first, count = shard_range(TOTAL, rank, world)
    env = OracleBatch(count, env_id_base=first, **KW)
    acts = np.random.RandomState(5).randint(0, 4, (T, TOTAL, KW["n_drones"])).astype(np.uint8)
    rew_sum = np.zeros(count)
    for t in range(T):
        _, r, _ = env.step(acts[t, first:first + count])
        rew_sum += r
    stats = torch.from_numpy(env.stats.copy())
    reduce_stats(stats)
    gathered = [None] * world
    dist.all_gather_object(gathered, (first, env.grid.copy(), env.pos.copy(), env.meta.copy(), rew_sum))
    if rank == 0:
        out.put((stats.numpy(), gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_partitions():
    for total in (0, 1, 7, 64, 65536 + 3):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (f0, c0), (f1, _) in zip(spans, spans[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


@pytest.mark.timeout(120)
def test_two_rank_gloo_matches_unsharded():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    stats, gathered = out.get(timeout=100)
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    ref = OracleBatch(TOTAL, env_id_base=0, **KW)
    acts = np.random.RandomState(5).randint(0, 4, (T, TOTAL, KW["n_drones"])).astype(np.uint8)
    rew_sum = np.zeros(TOTAL)
    for t in range(T):
        _, r, _ = ref.step(acts[t])
        rew_sum += r
    assert np.array_equal(stats, ref.stats)
    for first, grid, pos, meta, rs in gathered:
        n = grid.shape[0]
        assert np.array_equal(grid, ref.grid[first:first + n])
        assert np.array_equal(pos, ref.pos[first:first + n])
        assert np.array_equal(meta, ref.meta[first:first + n])
        assert np.array_equal(rs, rew_sum[first:first + n])
