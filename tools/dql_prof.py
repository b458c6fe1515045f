"""usage (GPU box): python tools/dql_prof.py [N] [D] [grid] [iters] — wall/device time of BatchedDQLAgent.train_step at the
cfg4 env shape (16384 envs x 8 drones, 10x10), printed as one JSON line (agent-steps/s, env step + act + replay)."""
import json
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(REPO, "marl-drones_b200"), REPO]
import torch  # noqa: E402

from swarm_b200 import BatchedDroneEnv  # noqa: E402
from swarm_b200.dql import BatchedDQLAgent  # noqa: E402

N, D, G, K = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((1, 16384), (2, 8), (3, 10), (4, 50)))
env = BatchedDroneEnv(N, row_count=G, col_count=G, n_drones=D, step_size=1.0, reward_mode="DQL", obs_mode="full",
                      auto_reset=True, seed=1)
agent = BatchedDQLAgent(env, memory_steps=48, batch_size=32 * N, seed=1)
for _ in range(40):                                                         # fill the memory past batch_size, warm up replay
    agent.train_step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(K):
    _, _, loss = agent.train_step()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
print(json.dumps({"workload": "cfg4_dql", "envs": N, "n_drones": D, "grid": G, "replay_batch": agent.batch_size,
                  "ms_per_train_step": ms, "wall_ms_per_train_step": (time.perf_counter() - t0) * 1e3 / K,
                  "agent_steps_per_s": N * D / (ms * 1e-3), "loss": float(loss), "epsilon": agent.epsilon}))
