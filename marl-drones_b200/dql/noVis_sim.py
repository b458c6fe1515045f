"""Drop-in module for `from noVis_sim import DroneEnv` in DQL_model/DQLearning.py:4 and DQL_model/Agent_test.py:7:
the DQL reward tail (DQL_model/noVis_sim.py:138-148)."""
from swarm_b200.facade import DQLDroneEnv as DroneEnv  # noqa: F401
