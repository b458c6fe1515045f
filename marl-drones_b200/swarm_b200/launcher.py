"""Runs an UNMODIFIED reference script against the GPU env: pre-seeds sys.modules['noVis_sim'] with the facade module
and then executes the script as __main__ (the scripts resolve `noVis_sim` from their own directory first, so a
PYTHONPATH entry alone would not win).

    python -m swarm_b200.launcher --variant DQL /path/to/MARL-Drones/DQL_model/DQLearning.py --episodes 1
"""
import argparse
import importlib.util
import os
import runpy
import sys

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def facade_module(variant="A2C"):
    path = os.path.join(_PKG_ROOT, "noVis_sim.py" if variant == "A2C" else os.path.join("dql", "noVis_sim.py"))
    spec = importlib.util.spec_from_file_location("noVis_sim", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def install(variant="A2C"):
    """sys.modules['noVis_sim'] = facade; returns the module."""
    if _PKG_ROOT not in sys.path:
        sys.path.insert(0, _PKG_ROOT)
    mod = facade_module(variant)
    sys.modules["noVis_sim"] = mod
    return mod


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--variant", choices=["A2C", "DQL"], default="A2C")
    ap.add_argument("script")
    ap.add_argument("args", nargs=argparse.REMAINDER)
    ns = ap.parse_args(argv)
    install(ns.variant)
    sys.argv = [ns.script] + ns.args
    sys.path.insert(0, os.path.dirname(os.path.abspath(ns.script)))
    runpy.run_path(ns.script, run_name="__main__")


if __name__ == "__main__":
    main()
