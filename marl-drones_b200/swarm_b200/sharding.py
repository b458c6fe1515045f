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


class StatsComm:
    """The library's own NCCL communicator for the statistics all-reduce (C ABI: swarm_comm_* / swarm_stats_allreduce).

    Rank 0 draws the ncclUniqueId, the 128 bytes travel through the caller's torch.distributed group (any backend), every
    rank initialises the communicator on its current CUDA device.  `allreduce(stats)` enqueues ncclAllReduce(SUM, int64[8])
    on the current stream — asynchronous and CUDA-graph capturable, no torch.distributed call on the hot loop.
    """

    def __init__(self, group=None):
        import ctypes as C
        import torch.distributed as dist
        from . import _lib
        self._L = _lib.lib()
        on = dist.is_available() and dist.is_initialized()
        self.rank, self.world = (dist.get_rank(group), dist.get_world_size(group)) if on else (0, 1)
        uid = (C.c_uint8 * 128)()
        if self.rank == 0:
            _lib.check(self._L.swarm_comm_unique_id(uid))
        if self.world > 1:
            box = [bytes(uid)]
            dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            uid = (C.c_uint8 * 128).from_buffer_copy(box[0])
        comm = C.c_void_p()
        _lib.check(self._L.swarm_comm_init(uid, self.rank, self.world, C.byref(comm)))
        self._comm = comm

    def allreduce(self, stats, env=None):
        """In-place SUM over ranks of a device int64[8] tensor (or, with stats=None, of `env`'s internal statistics)."""
        import ctypes as C
        from . import _lib
        if stats is not None and (stats.dtype != torch.int64 or stats.numel() != 8 or not stats.is_cuda or not stats.is_contiguous()):
            raise ValueError("stats must be a contiguous CUDA int64[8] tensor")
        dev = stats.device if stats is not None else env.device
        _lib.check(self._L.swarm_stats_allreduce(env._h if env is not None else None, None if stats is None else C.c_void_p(stats.data_ptr()),
                                                 self._comm, C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
        return stats

    def close(self):
        if getattr(self, "_comm", None) is not None and self._comm.value:
            self._L.swarm_comm_destroy(self._comm)
            self._comm = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_sharded_env(total_envs, rank=None, world=None, **kw):
    """BatchedDroneEnv over this rank's shard of `total_envs` global envs (rank/world default to torch.distributed's)."""
    import torch.distributed as dist
    from .batched import BatchedDroneEnv
    if rank is None or world is None:
        on = dist.is_available() and dist.is_initialized()
        rank, world = (dist.get_rank(), dist.get_world_size()) if on else (0, 1)
    first, count = shard_range(total_envs, rank, world)
    return BatchedDroneEnv(count, env_id_base=kw.pop("env_id_base", 0) + first, **kw)
