"""Single-env facade with the reference's exact surface: `DroneEnv(...)`, `reset() -> int64[R*C]`,
`step(actions) -> (obs, reward, done)` (3-tuple, A2C/noVis_sim.py:175) and the attributes the reference's callers read
(`action_size`, `state_size`, `n_drones`, `n_drones_pos`, `uncertain_points`, `grid`, `step_func_count`;
A2C/main.py:105-107, A2C/train.py:54,81,229,266, A2C/a2c_eval.py:77,92,97, DQL_model/Agent_test.py:168).

The env itself lives on the GPU (a BatchedDroneEnv of one env, auto_reset off); every step is ONE C-ABI call
(`swarm_step_host_pos`: one kernel launch that reads the pinned action buffer and writes obs / reward / done into pinned host
memory, the position copy, one synchronisation).  `reset()` draws from the GLOBAL legacy `np.random` stream in the reference's order
(A2C/noVis_sim.py:47-49,59-61), so `np.random.seed(s)` reproduces the reference's resets bit for bit.
"""
import ctypes as C

import numpy as np
import torch

from ._lib import check
from .batched import BatchedDroneEnv, _raw_stream


class DroneEnv:
    reward_mode = "A2C"   # DQLDroneEnv overrides (DQL_model/noVis_sim.py:138-148)

    def __init__(self, row_count=10, col_count=10, n_anamolous=5, uncertainity=5, collision_dist=0.5, n_drones=3,
                 step_size=0.5):
        self.row_count, self.col_count = row_count, col_count
        self.n_anamolous, self.uncertainity = n_anamolous, uncertainity
        self.collision_dist, self.n_drones, self.step_size = collision_dist, n_drones, step_size
        self.action_size = 4
        self.state_size = row_count * col_count
        self.step_func_count = 0
        self._env = BatchedDroneEnv(1, row_count, col_count, n_anamolous, uncertainity, collision_dist, n_drones, step_size,
                                    reward_mode=self.reward_mode, obs_mode="full", obs_dtype=torch.uint8, auto_reset=False)
        # pinned host buffers of ONE swarm_step_host_pos call: at this size the kernel reads the actions and writes obs / reward /
        # done straight over PCIe, so a step is one launch, the position copy and one synchronisation
        pin = lambda shape, dt: torch.zeros(shape, dtype=dt).pin_memory()
        self._act_host, self._obs_host = pin((1, n_drones), torch.uint8), pin(self.state_size, torch.uint8)
        self._rew_host, self._done_host = pin(1, torch.float32), pin(1, torch.uint8)
        self._pos_host = pin((n_drones, 2), torch.float32)
        self._act_np, self._obs_np, self._pos_np = self._act_host.numpy(), self._obs_host.numpy(), self._pos_host.numpy()
        self._rew_np, self._done_np = self._rew_host.numpy(), self._done_host.numpy()
        fn, h = self._env._L.swarm_step_host_pos, self._env._h
        ptrs = [C.c_void_p(t.data_ptr()) for t in (self._act_host, self._obs_host, self._rew_host, self._done_host, self._pos_host)]
        dev_index = self._env.device.index
        def step_call():
            rc = fn(h, ptrs[0], ptrs[1], ptrs[2], ptrs[3], ptrs[4], None, _raw_stream(dev_index))
            if rc:
                check(rc)
        self._step_call = step_call
        self._obs_u8 = np.ones(self.state_size, np.uint8)
        self.n_drones_pos = None
        self.uncertain_points = None
        self.init_env()

    # ---------------------------------------------------------------------------------------------- reference API
    def init_env(self):
        """A2C/noVis_sim.py:36-63, same np.random draw order; the state is then injected into the device env."""
        R, Cc = self.row_count, self.col_count
        self.step_func_count = 0
        grid = np.ones((R, Cc), np.int8)
        self._anom = []
        placed = 0
        while placed < self.n_anamolous:
            a = np.random.randint(R)
            b = np.random.randint(Cc)
            if grid[a, b] == 1:
                grid[a, b] = self.uncertainity
                self._anom.append((a, b))
                placed += 1
        pos = np.zeros((self.n_drones, 2), np.float32)
        for i in range(self.n_drones):
            pos[i, 0] = np.random.randint(R)
            pos[i, 1] = np.random.randint(Cc)
        grid_t = np.ascontiguousarray(grid.T)
        self._env.set_state(grid_t[None], pos[None], np.zeros(1, np.int32))
        self._obs_u8 = grid_t.reshape(-1).astype(np.uint8)
        self._set_pos(pos)
        self.uncertain_points = {k: 1 for k in self._anom}

    def reset(self):
        self.init_env()
        return self._obs_u8.astype(np.int64)

    def step(self, actions):
        assert len(actions) == self.n_drones, "provide actions for each drone"
        host = self._act_np
        for i, action in enumerate(actions):
            assert action >= 0 and action < self.action_size, f"action should be in range:[0,{self.action_size})"
            host[0, i] = action if type(action) is int else int(np.asarray(action).reshape(-1)[0])
        self._step_call()                       # H2D (or direct read), the fused step, obs / reward / done / positions back, sync
        self._obs_u8 = self._obs_np.copy()
        reward = float(self._rew_np[0])
        done_b = bool(self._done_np[0])
        self._set_pos(self._pos_np)
        self.step_func_count += 1
        for k in self.uncertain_points:
            if self.uncertain_points[k] == 1 and self._obs_u8[k[1] * self.row_count + k[0]] == 0:
                self.uncertain_points[k] = 0
        return self._obs_u8.astype(np.int64), reward, done_b

    # ---------------------------------------------------------------------------------------------- attributes
    def _set_pos(self, pos):
        self.n_drones_pos = [[np.float64(p[0]), np.float64(p[1])] for p in pos]

    @property
    def grid(self):
        """reference `self.grid` (list of R lists of C ints, grid[x][y])."""
        return self._obs_u8.reshape(self.col_count, self.row_count).T.astype(np.int64).tolist()

    def close(self):
        self._env.close()


class DQLDroneEnv(DroneEnv):
    reward_mode = "DQL"
