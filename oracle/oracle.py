"""ctypes loader for the CPU oracle (oracle/swarm_oracle.c) — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
The product path (marl-drones_b200/) never does; it has no CPU fallback.

`OracleBatch` keeps N envs as numpy arrays in the same layouts the C ABI exposes
(include/swarm_b200.h): grid int8 [N, C, R] (== reference grid[x][y] transposed, i.e. the obs order of
A2C/noVis_sim.py:168-170), pos float64 [N, D, 2], meta int32 [N, 4], wait int16 [N, C, R].
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_swarm.so")


class OracleConfig(C.Structure):
    _fields_ = [
        ("row_count", C.c_int32), ("col_count", C.c_int32), ("n_drones", C.c_int32),
        ("n_anamolous", C.c_int32), ("uncertainity", C.c_int32),
        ("reward_mode", C.c_int32), ("obs_mode", C.c_int32), ("window", C.c_int32),
        ("auto_reset", C.c_int32), ("dwell_normal", C.c_int32), ("dwell_anom", C.c_int32),
        ("scan_mode", C.c_int32),
        ("collision_dist", C.c_double), ("step_size", C.c_double),
        ("seed", C.c_uint64), ("env_id_base", C.c_int64),
    ]


def build(force=False):
    """Compile the oracle with gcc (oracle/Makefile). Building the checker is not using it."""
    src = os.path.join(_HERE, "swarm_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle_swarm.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        vp = C.c_void_p
        L.oracle_step_batch.argtypes = [C.POINTER(OracleConfig), C.c_int32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int32]
        L.oracle_step_batch.restype = None
        L.oracle_reset_batch.argtypes = [C.POINTER(OracleConfig), C.c_int32, vp, vp, vp, vp, vp, vp]
        L.oracle_reset_batch.restype = None
        L.oracle_init_batch.argtypes = [C.POINTER(OracleConfig), C.c_int32, vp, vp, vp, vp]
        L.oracle_init_batch.restype = None
        L.oracle_emit_obs_batch.argtypes = [C.POINTER(OracleConfig), C.c_int32, vp, vp, vp]
        L.oracle_emit_obs_batch.restype = None
        L.oracle_reset_numpy.argtypes = [C.POINTER(OracleConfig), C.c_uint32, C.c_int32, vp, vp, vp, vp]
        L.oracle_reset_numpy.restype = None
        L.oracle_greedy_actions.argtypes = [C.POINTER(OracleConfig), C.c_int32, vp, vp, vp]
        L.oracle_greedy_actions.restype = None
        L.oracle_drone_update.argtypes = [C.POINTER(OracleConfig), C.c_int32, vp, vp, vp, vp]
        L.oracle_drone_update.restype = None
        L.oracle_philox4x32_10.argtypes = [vp, vp, vp]
        L.oracle_philox4x32_10.restype = None
        L.oracle_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleBatch:
    """N independent reference-semantics envs on the CPU."""

    def __init__(self, n_envs=1, row_count=10, col_count=10, n_anamolous=5, uncertainity=5, collision_dist=0.5,
                 n_drones=3, step_size=0.5, reward_mode=0, obs_mode=0, window=5, auto_reset=False,
                 dwell_normal=0, dwell_anom=0, seed=0, env_id_base=0, scan_mode=0, nthreads=1):
        self.N, self.R, self.Cc, self.D = n_envs, row_count, col_count, n_drones
        self.cfg = OracleConfig(row_count, col_count, n_drones, n_anamolous, uncertainity, reward_mode, obs_mode,
                                window, int(auto_reset), dwell_normal, dwell_anom, scan_mode,
                                float(collision_dist), float(step_size), seed, env_id_base)
        self.nthreads = nthreads
        self.grid = np.ones((n_envs, col_count, row_count), np.int8)
        self.pos = np.zeros((n_envs, n_drones, 2), np.float64)
        self.meta = np.zeros((n_envs, 4), np.int32)
        self.wait = np.zeros((n_envs, col_count, row_count), np.int16) if (dwell_normal or dwell_anom) else None
        self.stats = np.zeros(8, np.int64)
        self.bad_action = np.zeros(1, np.int32)
        self.obs_per_env = row_count * col_count if obs_mode == 0 else n_drones * window * window
        lib().oracle_init_batch(C.byref(self.cfg), self.N, _p(self.grid), _p(self.pos), _p(self.meta), _p(self.wait))

    # -- state -------------------------------------------------------------------------------------------------
    def set_state(self, grid, pos, t, wait=None, first=0):
        """grid [n, C, R] (transposed/obs order), pos [n, D, 2], t [n]."""
        n = grid.shape[0]
        self.grid[first:first + n] = grid
        self.pos[first:first + n] = pos
        self.meta[first:first + n, 0] = t
        self.meta[first:first + n, 1] = (np.asarray(grid).reshape(n, -1) > 0).sum(1)
        self.meta[first:first + n, 2] = 0
        if self.wait is not None and wait is not None:
            self.wait[first:first + n] = wait

    def reset_numpy(self, env, seed, n_resets=1):
        """State of `np.random.seed(seed)` followed by n_resets init_env() calls (A2C/noVis_sim.py:36-63)."""
        lib().oracle_reset_numpy(C.byref(self.cfg), seed, n_resets, _p(self.grid[env]), _p(self.pos[env]),
                                 _p(self.meta[env]), _p(self.wait[env]) if self.wait is not None else None)

    # -- API ---------------------------------------------------------------------------------------------------
    def obs(self):
        out = np.empty((self.N, self.obs_per_env), np.float32)
        lib().oracle_emit_obs_batch(C.byref(self.cfg), self.N, _p(self.grid), _p(self.pos), _p(out))
        return out

    def reset(self, mask=None):
        out = np.empty((self.N, self.obs_per_env), np.float32)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        lib().oracle_reset_batch(C.byref(self.cfg), self.N, _p(m), _p(self.grid), _p(self.pos), _p(self.meta),
                                 _p(self.wait), _p(out))
        return out

    def step(self, actions, want_obs=True):
        a = np.ascontiguousarray(actions, np.uint8).reshape(self.N, self.D)
        obs = np.empty((self.N, self.obs_per_env), np.float32) if want_obs else None
        rew = np.empty(self.N, np.float32)
        done = np.empty(self.N, np.uint8)
        lib().oracle_step_batch(C.byref(self.cfg), self.N, _p(self.grid), _p(self.pos), _p(self.meta), _p(self.wait),
                                _p(a), _p(obs), _p(rew), _p(done), _p(self.stats), _p(self.bad_action), self.nthreads)
        return obs, rew, done

    def greedy_actions(self):
        out = np.empty((self.N, self.D), np.uint8)
        lib().oracle_greedy_actions(C.byref(self.cfg), self.N, _p(self.grid), _p(self.pos), _p(out))
        return out


    def drone_update(self):
        """(target int32 [N, D] flat cell or -1, dir int8 [N, D, 2]) of swarmsim-2d/drone.py:29-59 on the lattice."""
        target = np.empty((self.N, self.D), np.int32)
        direction = np.empty((self.N, self.D, 2), np.int8)
        lib().oracle_drone_update(C.byref(self.cfg), self.N, _p(self.grid), _p(self.pos), _p(target), _p(direction))
        return target, direction


def philox(ctr, key):
    c = np.asarray(ctr, np.uint32)
    k = np.asarray(key, np.uint32)
    out = np.zeros(4, np.uint32)
    lib().oracle_philox4x32_10(_p(c), _p(k), _p(out))
    return out
