"""ctypes binding of the C ABI in include/swarm_b200.h (libswarm_b200.so, built in-tree by build()).

There is no CPU fallback: if the library is missing it is (re)built with nvcc, and if it cannot be loaded or no
sm_100 device exists every entry point raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG_ROOT = os.path.dirname(_HERE)                      # marl-drones_b200/
_REPO = os.path.dirname(_PKG_ROOT)
CSRC = os.path.join(_PKG_ROOT, "csrc")
SO_PATH = os.environ.get("SWARM_B200_SO") or os.path.join(_HERE, "libswarm_b200.so")   # override: kernel-variant experiments
HEADER = os.path.join(_REPO, "include", "swarm_b200.h")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-ldl"]

REWARD_A2C, REWARD_DQL = 0, 1
OBS_FULL, OBS_WINDOW = 0, 1
DTYPE_F32, DTYPE_U8, DTYPE_P2 = 0, 1, 2
ACT_U8, ACT_I32, ACT_I64 = 0, 1, 2
STAT_NAMES = ("env_steps", "agent_steps", "reward_sum", "episodes_done", "episode_len_sum", "episode_return_sum",
              "cells_mapped", "anomalous_mapped")


class SwarmConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("n_envs", C.c_int32), ("n_drones", C.c_int32), ("row_count", C.c_int32),
        ("col_count", C.c_int32), ("n_anamolous", C.c_int32), ("uncertainity", C.c_int32), ("reward_mode", C.c_int32),
        ("obs_mode", C.c_int32), ("window", C.c_int32), ("obs_dtype", C.c_int32), ("auto_reset", C.c_int32),
        ("dwell_normal", C.c_int32), ("dwell_anom", C.c_int32), ("device", C.c_int32), ("reserved0", C.c_int32),
        ("collision_dist", C.c_double), ("step_size", C.c_double), ("seed", C.c_uint64), ("env_id_base", C.c_int64),
    ]


class SwarmError(RuntimeError):
    pass


def sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh"))] + [HEADER]


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> swarm_b200/libswarm_b200.so (in-tree, travels with gpurun)."""
    srcs = sources()
    if not force and os.path.exists(SO_PATH) and all(os.path.getmtime(SO_PATH) >= os.path.getmtime(s) for s in srcs):
        return SO_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(nvcc):
        nvcc = "nvcc"
    cmd = [nvcc] + NVCC_FLAGS + os.environ.get("SWARM_NVCC_EXTRA", "").split() + ["-o", SO_PATH, os.path.join(CSRC, "swarm_api.cu")]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return SO_PATH


_lib = None


def lib():
    """Loads libswarm_b200.so (building it if the sources are newer). Raises if that is impossible."""
    global _lib
    if _lib is not None:
        return _lib
    try:
        build()
    except Exception as exc:   # a stale-but-present library is still usable on a box without nvcc
        if not os.path.exists(SO_PATH):
            raise SwarmError(f"libswarm_b200.so is missing and could not be built: {exc}") from exc
    L = C.CDLL(SO_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    cfgp = C.POINTER(SwarmConfig)
    sig = {
        "swarm_abi_version": (C.c_int, []),
        "swarm_strerror": (C.c_char_p, [C.c_int]),
        "swarm_last_cuda_error": (C.c_char_p, []),
        "swarm_device_count": (C.c_int, []),
        "swarm_create": (C.c_int, [cfgp, C.POINTER(vp)]),
        "swarm_destroy": (C.c_int, [vp]),
        "swarm_get_config": (C.c_int, [vp, cfgp]),
        "swarm_obs_elems": (i64, [vp]),
        "swarm_obs_bytes": (i64, [vp]),
        "swarm_state_ptr": (C.c_int, [vp, C.c_int, C.POINTER(vp), C.POINTER(i64)]),
        "swarm_reset": (C.c_int, [vp, vp, vp, vp]),
        "swarm_set_state": (C.c_int, [vp, vp, vp, vp, vp, i32, i32, vp]),
        "swarm_get_state": (C.c_int, [vp, vp, vp, vp, vp, i32, i32, vp]),
        "swarm_step": (C.c_int, [vp, vp, C.c_int, vp, vp, vp, vp, vp]),
        "swarm_rollout": (C.c_int, [vp, i32, vp, vp, i32, vp, vp, vp, vp]),
        "swarm_step_host": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
        "swarm_step_host_static": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
        "swarm_step_host_pos": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, vp]),
        "swarm_stage_move": (C.c_int, [vp, vp, C.c_int, vp]),
        "swarm_stage_scatter": (C.c_int, [vp, vp, vp]),
        "swarm_stage_reward": (C.c_int, [vp, vp, vp, vp, vp]),
        "swarm_stage_obs": (C.c_int, [vp, vp, vp]),
        "swarm_greedy_actions": (C.c_int, [vp, vp, vp]),
        "swarm_device_error": (C.c_int, [vp, C.POINTER(i32), C.c_int, vp]),
        "swarm_launch_count": (i64, [vp]),
        "swarm_comm_unique_id": (C.c_int, [vp]),
        "swarm_comm_init": (C.c_int, [vp, i32, i32, C.POINTER(vp)]),
        "swarm_comm_destroy": (C.c_int, [vp]),
        "swarm_stats_allreduce": (C.c_int, [vp, vp, vp, vp]),
        "swarm_stats_mailbox_handle": (C.c_int, [vp, vp]),
        "swarm_stats_peers_open": (C.c_int, [vp, vp, i32, i32]),
        "swarm_stats_peers_close": (C.c_int, [vp]),
        "swarm_stats_global": (C.c_int, [vp, i64, vp, vp]),
        "swarm_stats_push": (C.c_int, [vp, vp, vp]),
        "swarm_stats_global_host": (C.c_int, [vp, i64, vp, vp]),
        "swarm_sample_actions": (C.c_int, [vp, i32, i64, i32, C.c_uint64, C.c_uint32, i64, vp, vp]),
        "swarm_nstep_returns": (C.c_int, [vp, vp, vp, i32, i64, i32, C.c_float, vp, vp, vp]),
        "swarm_gru_gates_fwd": (C.c_int, [vp, vp, i64, i32, i32, vp, vp]),
        "swarm_bias_act": (C.c_int, [vp, vp, i32, i64, i32, i32, vp]),
        "swarm_embed_tanh_l2norm_fwd": (C.c_int, [vp, vp, vp, vp, i64, i32, i32, i32, C.c_float, vp, vp]),
        "swarm_embed_tanh_l2norm_bwd": (C.c_int, [vp, vp, vp, vp, vp, i64, i32, i32, i32, C.c_float, vp, vp, vp]),
        "swarm_row_standardize": (C.c_int, [vp, i64, i32, vp, vp]),
        "swarm_relu_l2norm_fwd": (C.c_int, [vp, i64, i32, C.c_float, vp, vp]),
        "swarm_relu_l2norm_bwd": (C.c_int, [vp, vp, i64, i32, C.c_float, vp, vp]),
        "swarm_gru_gates_bwd": (C.c_int, [vp, vp, vp, i64, i32, i32, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    if L.swarm_abi_version() != 1:
        raise SwarmError("libswarm_b200.so ABI version mismatch")
    _lib = L
    return L


EXPORTED = ("swarm_abi_version swarm_strerror swarm_last_cuda_error swarm_device_count swarm_create swarm_destroy "
            "swarm_get_config swarm_obs_elems swarm_obs_bytes swarm_state_ptr swarm_reset swarm_set_state swarm_get_state "
            "swarm_step swarm_rollout swarm_step_host swarm_stage_move swarm_stage_scatter swarm_stage_reward "
            "swarm_stage_obs swarm_greedy_actions swarm_device_error swarm_launch_count swarm_sample_actions "
            "swarm_nstep_returns swarm_gru_gates_fwd swarm_gru_gates_bwd swarm_bias_act swarm_embed_tanh_l2norm_fwd swarm_embed_tanh_l2norm_bwd swarm_row_standardize swarm_relu_l2norm_fwd swarm_relu_l2norm_bwd swarm_comm_unique_id swarm_comm_init swarm_comm_destroy swarm_stats_allreduce swarm_stats_mailbox_handle "
            "swarm_stats_peers_open swarm_stats_peers_close swarm_stats_global swarm_stats_global_host swarm_stats_push "
            "swarm_step_host_static swarm_step_host_pos").split()


def check(status):
    if status != 0:
        L = lib()
        msg = L.swarm_strerror(status).decode()
        if status in (-3, -6, -7):
            msg += ": " + L.swarm_last_cuda_error().decode()
        raise SwarmError(f"swarm_b200 error {status}: {msg}")
