"""BatchedDroneEnv — the VecEnv-style surface (reference: A2C/envs.py:58-117, RenderSubprocVecEnv) over N
GPU-resident envs: `reset() -> obs[N, ...]`, `step(actions[N, D]) -> (obs, rewards[N], dones[N], infos)`,
`num_envs`, `close()`, auto-reset on done returning the reset obs (A2C/envs.py:22-26).

PyTorch only allocates the I/O tensors and provides the stream; every computation is a kernel of
libswarm_b200.so reached through the C ABI (include/swarm_b200.h).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import (ACT_I32, ACT_I64, ACT_U8, DTYPE_F32, DTYPE_P2, DTYPE_U8, OBS_FULL, OBS_WINDOW, REWARD_A2C, REWARD_DQL,
                   STAT_NAMES, SwarmConfig, SwarmError, check)

_ACT_DTYPES = {torch.uint8: ACT_U8, torch.int32: ACT_I32, torch.int64: ACT_I64}
_raw_stream = torch._C._cuda_getCurrentRawStream   # (device_index) -> cudaStream_t as int; far cheaper than current_stream()


class _DeviceView:
    """Zero-copy torch view of library-owned device memory (via __cuda_array_interface__)."""

    def __init__(self, ptr, shape, typestr, owner):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 2}
        self._owner = owner


class BatchedDroneEnv:
    def __init__(self, n_envs, row_count=10, col_count=10, n_anamolous=5, uncertainity=5, collision_dist=0.5,
                 n_drones=3, step_size=0.5, reward_mode="A2C", obs_mode="full", window=5, obs_dtype=torch.float32,
                 auto_reset=True, dwell_normal=0, dwell_anom=0, seed=0, env_id_base=0, device=None):
        if not torch.cuda.is_available():
            raise SwarmError("BatchedDroneEnv needs a CUDA device (sm_100); there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else torch.device(device).index or 0)
        self._L = _lib.lib()
        rm = {"A2C": REWARD_A2C, "DQL": REWARD_DQL}[reward_mode] if isinstance(reward_mode, str) else int(reward_mode)
        om = {"full": OBS_FULL, "window": OBS_WINDOW}[obs_mode] if isinstance(obs_mode, str) else int(obs_mode)
        if obs_dtype == "packed2":      # SWARM_DTYPE_P2: one int64 word per drone (5x5 windows as two bit planes; unpack_windows)
            obs_dtype = torch.int64
        od = {torch.float32: DTYPE_F32, torch.uint8: DTYPE_U8, torch.int64: DTYPE_P2}[obs_dtype]
        self.cfg = SwarmConfig(C.sizeof(SwarmConfig), n_envs, n_drones, row_count, col_count, n_anamolous, uncertainity,
                               rm, om, window, od, int(bool(auto_reset)), dwell_normal, dwell_anom, self.device.index, 0,
                               float(collision_dist), float(step_size), seed, env_id_base)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(self._L.swarm_create(C.byref(self.cfg), C.byref(h)))
        self._h = h
        self.num_envs = n_envs
        self.n_drones = n_drones
        self.row_count, self.col_count = row_count, col_count
        self.action_size = 4                                   # A2C/noVis_sim.py:29
        self.state_size = row_count * col_count                # A2C/noVis_sim.py:30
        self.obs_dtype = obs_dtype
        self.obs_shape = (n_envs, row_count * col_count) if om == OBS_FULL else (n_envs, n_drones, window * window)
        if od == DTYPE_P2:
            self.obs_shape = (n_envs, n_drones)
        self.has_wait = bool(dwell_normal or dwell_anom)
        self.stats = torch.zeros(8, dtype=torch.int64, device=self.device)
        self._obs = torch.empty(self.obs_shape, dtype=obs_dtype, device=self.device)
        self._rew = torch.empty(n_envs, dtype=torch.float32, device=self.device)
        self._done = torch.empty(n_envs, dtype=torch.uint8, device=self.device)
        # cached raw pointers / bound function for the per-step fast path (the tensors above live as long as self)
        self._nact = n_envs * n_drones
        self._dev_index = self.device.index
        self._obs_ptr, self._rew_ptr, self._done_ptr = self._obs.data_ptr(), self._rew.data_ptr(), self._done.data_ptr()
        self._stats_ptr = self.stats.data_ptr()
        self._info = {"stats": self.stats}
        self._step_fn = self._L.swarm_step
        ptr, nbytes = C.c_void_p(), C.c_int64()
        check(self._L.swarm_state_ptr(self._h, 0, C.byref(ptr), C.byref(nbytes)))
        #: resident positions float32 [N, D, 2] — a VIEW of the state (np.array(env.n_drones_pos), A2C/train.py:54,81)
        self.pos = torch.as_tensor(_DeviceView(ptr.value, (n_envs, n_drones, 2), "<f4", self), device=self.device)
        check(self._L.swarm_state_ptr(self._h, 1, C.byref(ptr), C.byref(nbytes)))
        #: resident meta int32 [N, 4] = {step_func_count, remaining, episode_return, episode} (view)
        self.meta = torch.as_tensor(_DeviceView(ptr.value, (n_envs, 4), "<i4", self), device=self.device)

    # ------------------------------------------------------------------------------------------------ helpers
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _chk_tensor(self, t, dtype, numel, name):
        if t.device != self.device or not t.is_contiguous() or t.dtype != dtype or t.numel() != numel:
            raise ValueError(f"{name}: expected contiguous {dtype} tensor with {numel} elements on {self.device}")

    @staticmethod
    def _ptr(t):
        return None if t is None else C.c_void_p(t.data_ptr())

    # ------------------------------------------------------------------------------------------------ VecEnv surface
    def reset(self, mask=None, out=None):
        """A2C/envs.py:95-98. mask: optional bool/uint8 [N] (only those envs restart). Returns obs of all envs."""
        obs = self._obs if out is None else out
        self._chk_tensor(obs, self.obs_dtype, self._obs.numel(), "out")
        m = None
        if mask is not None:
            m = mask.to(device=self.device, dtype=torch.uint8).contiguous()
        check(self._L.swarm_reset(self._h, self._ptr(m), self._ptr(obs), self._stream()))
        return obs

    def step(self, actions, out=None, want_obs=True):
        """A2C/envs.py:81-93. actions: integer tensor [N, D] on the device. Returns (obs, rewards, dones, infos);
        infos holds the int64[8] running statistics tensor, still on the device.
        The common case (contiguous uint8/int32/int64 CUDA tensor) goes straight to the C ABI: one kernel launch."""
        if not (torch.is_tensor(actions) and actions.dtype in _ACT_DTYPES and actions.device == self.device
                and actions.is_contiguous()):
            if not torch.is_tensor(actions):
                actions = torch.as_tensor(np.asarray(actions), device=self.device)
            if actions.dtype not in _ACT_DTYPES:
                actions = actions.to(torch.int64)
            actions = actions.to(self.device).contiguous()
        if actions.numel() != self._nact:
            raise AssertionError("provide actions for each drone")      # A2C/noVis_sim.py:114
        obs, optr = None, None
        if want_obs:
            if out is None:
                obs, optr = self._obs, self._obs_ptr
            else:
                self._chk_tensor(out, self.obs_dtype, self._obs.numel(), "out")
                obs, optr = out, out.data_ptr()
        rc = self._step_fn(self._h, actions.data_ptr(), _ACT_DTYPES[actions.dtype], optr, self._rew_ptr, self._done_ptr,
                           self._stats_ptr, _raw_stream(self._dev_index))
        if rc:
            check(rc)
        return obs, self._rew, self._done, self._info

    def rollout(self, actions, obs_out=None, obs_last_only=False, reward_out=None, done_out=None):
        """T steps in one launch. actions uint8 [T, N, D]. Returns (obs_out, rewards [T, N], dones [T, N])."""
        T = actions.shape[0]
        self._chk_tensor(actions, torch.uint8, T * self.num_envs * self.n_drones, "actions")
        if reward_out is None:
            reward_out = torch.empty((T, self.num_envs), dtype=torch.float32, device=self.device)
        if done_out is None:
            done_out = torch.empty((T, self.num_envs), dtype=torch.uint8, device=self.device)
        if obs_out is not None:
            self._chk_tensor(obs_out, self.obs_dtype, (1 if obs_last_only else T) * self._obs.numel(), "obs_out")
        check(self._L.swarm_rollout(self._h, T, self._ptr(actions), self._ptr(obs_out), int(obs_last_only),
                                    self._ptr(reward_out), self._ptr(done_out), self._ptr(self.stats), self._stream()))
        return obs_out, reward_out, done_out

    def step_host(self, actions_host, obs_host, reward_host, done_host, stats_host=None):
        """Host-buffer step (pinned CPU tensors / numpy arrays): H2D actions, step, D2H results, synchronous."""
        def hp(a):
            if a is None:
                return None
            return C.c_void_p(a.data_ptr() if torch.is_tensor(a) else a.ctypes.data)
        check(self._L.swarm_step_host(self._h, hp(actions_host), hp(obs_host), hp(reward_host), hp(done_host),
                                      hp(stats_host), self._stream()))

    def step_host_pos(self, actions_host, obs_host, reward_host, done_host, pos_host, stats_host=None):
        """step_host that also returns the positions after the step (float32 [N, D, 2], the reference's `n_drones_pos`)."""
        def hp(a):
            if a is None:
                return None
            return C.c_void_p(a.data_ptr() if torch.is_tensor(a) else a.ctypes.data)
        check(self._L.swarm_step_host_pos(self._h, hp(actions_host), hp(obs_host), hp(reward_host), hp(done_host),
                                          hp(pos_host), hp(stats_host), self._stream()))

    def step_host_static(self, actions_host, obs_host, reward_host, done_host, stats_host=None):
        """step_host through swarm_step_host_static: the device-side staging is the library's fixed static array
        (SWARM_STATIC_SCRATCH_BYTES), so the call allocates nothing; sizes beyond it raise SwarmError.  No counterpart in
        the reference (added for callers that cannot call cudaMalloc); calls that overlap on the GPU race."""
        def hp(a):
            if a is None:
                return None
            return C.c_void_p(a.data_ptr() if torch.is_tensor(a) else a.ctypes.data)
        rc = self._L.swarm_step_host_static(self._h, hp(actions_host), hp(obs_host), hp(reward_host), hp(done_host),
                                            hp(stats_host), self._stream())
        if rc == -1:
            msg = self._L.swarm_last_cuda_error().decode()
            if "static scratch" in msg:
                raise SwarmError(f"swarm_b200 error -1: {msg}")
        check(rc)

    def close(self, collective=True):
        """Destroys the handle.  With peer statistics enabled across ranks (enable_peer_stats on a process group) the close is
        COLLECTIVE: every rank stops pushing and unmaps its peers' mailboxes, then all ranks meet at a barrier, and only then
        is this rank's mailbox freed — a peer must never write into freed memory."""
        if getattr(self, "_h", None) is not None and self._h.value:
            torch.cuda.synchronize(self.device)
            if getattr(self, "_peer_world", 1) > 1:
                self._L.swarm_stats_peers_close(self._h)
                self._peer_world = 1
                if collective:
                    import torch.distributed as dist
                    if dist.is_available() and dist.is_initialized():
                        dist.barrier(group=self._peer_group)
            self._L.swarm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close(collective=False)
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------------ state
    def set_state(self, grid=None, pos=None, t=None, wait=None, first=0):
        """grid int8 [n, C, R] (obs order: grid[n, y, x] = reference grid[x][y]), pos float32 [n, D, 2], t int32 [n]."""
        def dev(a, dt):
            return None if a is None else torch.as_tensor(np.asarray(a) if not torch.is_tensor(a) else a).to(self.device, dt).contiguous()
        g, p, tt, w = dev(grid, torch.int8), dev(pos, torch.float32), dev(t, torch.int32), dev(wait, torch.int16)
        count = next(x.shape[0] for x in (g, p, tt, w) if x is not None)
        check(self._L.swarm_set_state(self._h, self._ptr(g), self._ptr(p), self._ptr(tt), self._ptr(w), first, count, self._stream()))
        torch.cuda.current_stream(self.device).synchronize()   # the temporaries above must outlive the kernel

    def get_state(self, first=0, count=None):
        count = self.num_envs - first if count is None else count
        g = torch.empty((count, self.col_count, self.row_count), dtype=torch.int8, device=self.device)
        p = torch.empty((count, self.n_drones, 2), dtype=torch.float32, device=self.device)
        t = torch.empty(count, dtype=torch.int32, device=self.device)
        w = torch.empty((count, self.col_count, self.row_count), dtype=torch.int16, device=self.device) if self.has_wait else None
        check(self._L.swarm_get_state(self._h, self._ptr(g), self._ptr(p), self._ptr(t), self._ptr(w), first, count, self._stream()))
        return g, p, t, w

    # ------------------------------------------------------------------------------------------------ stages / misc
    def stage_move(self, actions):
        actions = actions.to(self.device).contiguous()
        check(self._L.swarm_stage_move(self._h, self._ptr(actions), _ACT_DTYPES[actions.dtype], self._stream()))

    def stage_scatter(self):
        cover = torch.empty(self.num_envs, dtype=torch.int32, device=self.device)
        check(self._L.swarm_stage_scatter(self._h, self._ptr(cover), self._stream()))
        return cover

    def stage_reward(self, cover):
        rew = torch.empty(self.num_envs, dtype=torch.float32, device=self.device)
        done = torch.empty(self.num_envs, dtype=torch.uint8, device=self.device)
        check(self._L.swarm_stage_reward(self._h, self._ptr(cover), self._ptr(rew), self._ptr(done), self._stream()))
        return rew, done

    def stage_obs(self, out=None):
        obs = torch.empty(self.obs_shape, dtype=self.obs_dtype, device=self.device) if out is None else out
        check(self._L.swarm_stage_obs(self._h, self._ptr(obs), self._stream()))
        return obs

    def greedy_actions(self):
        a = torch.empty((self.num_envs, self.n_drones), dtype=torch.uint8, device=self.device)
        check(self._L.swarm_greedy_actions(self._h, self._ptr(a), self._stream()))
        return a

    def device_error(self, clear=True):
        code = C.c_int32()
        check(self._L.swarm_device_error(self._h, C.byref(code), int(clear), self._stream()))
        return code.value

    @property
    def launch_count(self):
        return int(self._L.swarm_launch_count(self._h))

    def stats_dict(self):
        return dict(zip(STAT_NAMES, self.stats.tolist()))

    # -------------------------------------------------------------------------------------- peer-memory statistics
    def enable_peer_stats(self, group=None):
        """Multi-GPU (one process per GPU of one node): exchange the ranks' mailbox handles through torch.distributed once;
        from then on every step launch publishes this rank's running statistics into every rank's mailbox over NVLink
        (swarm_stats_peers_open; no collective on the step path).  With no process group it is a world of one."""
        import torch.distributed as dist
        on = dist.is_available() and dist.is_initialized()
        rank, world = (dist.get_rank(group), dist.get_world_size(group)) if on else (0, 1)
        h = (C.c_uint8 * 64)()
        rc = self._L.swarm_stats_mailbox_handle(self._h, h)
        mine = bytes(h) if rc == 0 else None       # a rank that cannot export still takes part in the gather (no deadlock)
        handles = [mine]
        if world > 1:
            handles = [None] * world
            dist.all_gather_object(handles, mine, group=group)
        if any(x is None for x in handles):
            check(rc)
            raise SwarmError("another rank could not export its statistics mailbox (CUDA IPC)")
        buf = (C.c_uint8 * (64 * world)).from_buffer_copy(b"".join(handles))
        check(self._L.swarm_stats_peers_open(self._h, buf, rank, world))
        self._peer_world, self._peer_group = world, group
        self._gstats = torch.zeros(8, dtype=torch.int64, device=self.device)
        return self

    def push_stats(self):
        """Publish the CURRENT running totals to every rank's mailbox (stream-ordered after the steps enqueued so far)."""
        check(self._L.swarm_stats_push(self._h, self._stats_ptr, self._stream()))

    def global_stats(self, min_pushes=0, out=None):
        """int64[8] (device): the sum over all ranks of their latest pushed running totals (waits until every rank has pushed
        at least `min_pushes` times).  Every step launch of a rank publishes, at its start, the totals as the previous launch
        left them (the feed lags by one step and costs the step nothing); `push_stats()` publishes the current totals."""
        out = self._gstats if out is None else out
        check(self._L.swarm_stats_global(self._h, int(min_pushes), self._ptr(out), self._stream()))
        return out

    def global_stats_host(self, min_pushes=0):
        """The same as a list of 8 ints on the host; synchronises, and raises SwarmError (SWARM_ERR_TIMEOUT) when some rank's
        slot never became consistent instead of returning a torn sum."""
        out = (C.c_int64 * 8)()
        check(self._L.swarm_stats_global_host(self._h, int(min_pushes), out, self._stream()))
        return list(out)

    def disable_peer_stats(self):
        check(self._L.swarm_stats_peers_close(self._h))
        self._peer_world = 1

    def allreduce_stats(self, group=None):
        """The only collective on this path: ncclAllReduce(SUM, int64[8]) of the episode statistics (64 B per rank)."""
        from .sharding import reduce_stats
        return reduce_stats(self.stats.clone(), group=group)


def unpack_windows(packed, uncertainity=5, dtype=torch.float32):
    """Expands packed 2-bit windows (obs_dtype="packed2", SWARM_DTYPE_P2: int64 [..., D]) into [..., D, 25] window values:
    bit k = LO plane, bit 32 + k = HI plane of cell k = b*5 + a; code LO + 2 HI: 0 -> 0, 1 -> 1, 3 -> uncertainity, 2 -> -1
    (255 for uint8).  Works on CPU and CUDA tensors (and numpy arrays)."""
    if not torch.is_tensor(packed):
        packed = torch.as_tensor(np.asarray(packed).astype(np.int64))
    k = torch.arange(25, device=packed.device, dtype=torch.int64)
    p = packed.to(torch.int64).unsqueeze(-1)
    code = ((p >> k) & 1) + 2 * ((p >> (k + 32)) & 1)
    table = torch.tensor([0, 1, 255 if dtype == torch.uint8 else -1, uncertainity], device=packed.device).to(dtype)
    return table[code]
