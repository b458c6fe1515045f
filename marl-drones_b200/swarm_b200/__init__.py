"""swarm_b200 — B200-native batched drone-swarm coverage environment (drop-in for MARL-Drones' noVis_sim.DroneEnv).

Public surface:
  BatchedDroneEnv   VecEnv-style 4-tuple surface on CUDA tensors (reference: A2C/envs.py:58-117)
  DroneEnv          single-env 3-tuple facade, A2C reward (reference: A2C/noVis_sim.py:7-175)
  DQLDroneEnv       same with the DQL reward tail (reference: DQL_model/noVis_sim.py:138-148)
  make_sharded_env  this rank's contiguous shard of a global env batch (no data-path collective); reduce_stats = the one
                    collective (int64[8] statistics all-reduce); StatsComm = the same through
                    the library's own NCCL communicator (swarm_stats_allreduce)
  unpack_windows    expands the packed 2-bit window format (BatchedDroneEnv(obs_dtype="packed2"), SWARM_DTYPE_P2) into values
  build()           compile csrc/ for sm_100a into swarm_b200/libswarm_b200.so
"""
from ._lib import SwarmConfig, SwarmError, build, lib  # noqa: F401
from .batched import BatchedDroneEnv, unpack_windows  # noqa: F401
from .facade import DQLDroneEnv, DroneEnv  # noqa: F401
from .sharding import StatsComm, make_sharded_env, reduce_stats, shard_range  # noqa: F401

__all__ = ["BatchedDroneEnv", "DroneEnv", "DQLDroneEnv", "SwarmConfig", "SwarmError", "build", "lib", "make_sharded_env",
           "reduce_stats", "shard_range", "StatsComm", "unpack_windows"]
