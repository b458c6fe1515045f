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


def ref_dql_namespace():
    """{'AgentModel', 'Agent', ...} of the UNMODIFIED DQL_model/DQLearning.py.  The file is a script (argparse + the whole
    training run at module level), so only its definitions — everything above the `env = DroneEnv(` line — are executed,
    with `noVis_sim` resolved to the reference's own DQL env and an empty command line."""
    import sys
    if "DQLns" not in _cache:
        path = os.path.join(REF_ROOT, "DQL_model", "DQLearning.py")
        src = open(path).read()
        head = src[:src.index("\nenv = DroneEnv(")]
        saved_mod, saved_argv = sys.modules.get("noVis_sim"), sys.argv
        sys.modules["noVis_sim"], sys.argv = ref_module("DQL"), ["DQLearning.py"]
        try:
            ns = {"__name__": "_ref_DQLearning"}
            exec(compile(head, path, "exec"), ns)
        finally:
            sys.argv = saved_argv
            if saved_mod is None:
                del sys.modules["noVis_sim"]
            else:
                sys.modules["noVis_sim"] = saved_mod
        _cache["DQLns"] = ns
    return _cache["DQLns"]
