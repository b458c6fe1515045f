"""Loads the UNMODIFIED reference DroneEnv (both variants) when /root/reference is present (this container only;
the GPU box has no /root/reference, so everything that needs the live reference is skipped there and the
committed fixtures under tests/golden/ stand in)."""
import importlib.util
import os

import numpy as np

REF_ROOT = "/root/reference"
_VARIANTS = {"A2C": "A2C/noVis_sim.py", "DQL": "DQL_model/noVis_sim.py"}
_cache = {}


def have_reference():
    return os.path.exists(os.path.join(REF_ROOT, _VARIANTS["A2C"]))


def ref_module(variant="A2C"):
    if variant not in _cache:
        path = os.path.join(REF_ROOT, _VARIANTS[variant])
        spec = importlib.util.spec_from_file_location(f"_ref_noVis_sim_{variant}", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _cache[variant] = mod
    return _cache[variant]


def make_ref_env(variant="A2C", seed=None, **kw):
    if seed is not None:
        np.random.seed(seed)
    return ref_module(variant).DroneEnv(**kw)


def export_state(env):
    """(grid_t int8 [C, R], pos float64 [D, 2], t) from a reference env; grid_t[y][x] = env.grid[x][y]."""
    g = np.array(env.grid, dtype=np.int8).T.copy()
    pos = np.array(env.n_drones_pos, dtype=np.float64).reshape(env.n_drones, 2)
    return g, pos, int(env.step_func_count)
