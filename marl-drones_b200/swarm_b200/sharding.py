"""Env sharding across the GPUs of one box: envs are independent (A2C/noVis_sim.py:112-175 touches only its
own grid / positions / counter), so rank g owns a contiguous range of global env ids and nothing on the data path crosses
NVLink.  Device resets are keyed by the GLOBAL env id (`env_id_base`), so any sharding yields identical trajectories.
The only collective is the int64[8] episode-statistics all-reduce (64 B per rank; NCCL on GPUs, gloo in the CPU tests).
"""
import torch


def shard_range(total_envs, rank, world):
    """[first, first+count) of global env ids owned by `rank`; the remainder goes to the lowest ranks."""
    if not (0 <= rank < world) or total_envs < 0:
        raise ValueError("bad rank/world/total_envs")
    base, extra = divmod(total_envs, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def reduce_stats(stats, group=None, async_op=False):
    """SUM all-reduce of the int64[8] statistics vector in place (integer sums: reduction-order independent).
    Returns the work handle when async_op, else the tensor."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return stats
    if stats.dtype != torch.int64 or stats.numel() != 8:
        raise ValueError("stats must be int64[8]")
    work = dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
    return work if async_op else stats


def make_sharded_env(total_envs, rank=None, world=None, **kw):
    """BatchedDroneEnv over this rank's shard of `total_envs` global envs (rank/world default to torch.distributed's)."""
    import torch.distributed as dist
    from .batched import BatchedDroneEnv
    if rank is None or world is None:
        on = dist.is_available() and dist.is_initialized()
        rank, world = (dist.get_rank(), dist.get_world_size()) if on else (0, 1)
    first, count = shard_range(total_envs, rank, world)
    return BatchedDroneEnv(count, env_id_base=kw.pop("env_id_base", 0) + first, **kw)
