"""Shared helpers for the parity tests: golden-trace loading and config plumbing."""
import glob
import json
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    d = {k: z[k] for k in z.files}
    d["config"] = json.loads(str(d["config"]))
    return d


def env_kwargs(cfg):
    """reference kwargs + reward mode from a golden config dict."""
    kw = {k: cfg[k] for k in ("row_count", "col_count", "n_anamolous", "uncertainity", "collision_dist", "n_drones", "step_size")}
    return kw, (0 if cfg["variant"] == "A2C" else 1)


# The BASELINE.json shapes, scaled-down N for oracle-speed parity runs.
SHAPES = {
    "cfg1": dict(row_count=10, col_count=10, n_drones=3, step_size=0.5),
    "cfg2": dict(row_count=32, col_count=32, n_drones=4, step_size=1.0),
    "cfg3": dict(row_count=64, col_count=64, n_drones=16, step_size=1.0, obs_mode=1, window=5),
    "cfg4": dict(row_count=10, col_count=10, n_drones=8, step_size=1.0),
    "cfg5a": dict(row_count=256, col_count=256, n_drones=64, step_size=1.0),
    "cfg5b": dict(row_count=256, col_count=256, n_drones=64, step_size=1.0, obs_mode=1, window=5),
}
