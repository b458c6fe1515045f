"""Drop-in module for `from noVis_sim import DroneEnv` (A2C/main.py:5, A2C/a2c_eval.py:4, A2C/policy_test.py:8):
put this directory on sys.path ahead of the reference's A2C/ directory, or pre-seed sys.modules['noVis_sim'] with it
(swarm_b200.launcher does that).  A2C reward shaping (A2C/noVis_sim.py:141-152,172-173)."""
from swarm_b200.facade import DroneEnv  # noqa: F401
