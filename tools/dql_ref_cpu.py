"""usage (build container only — needs /root/reference): python tools/dql_ref_cpu.py [n_drones] [grid] [steps]
Times the UNMODIFIED reference deep-Q loop (DQL_model/DQLearning.py: Agent.act -> env.step -> memorize -> replay(32)) on the host,
driving its own classes exactly as its module-level loop does (:161-187), and prints one JSON line (agent-steps/s)."""
import json
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(REPO, "tests")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import refload  # noqa: E402

D, G, T = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((1, 8), (2, 10), (3, 120)))
ns = refload.ref_dql_namespace()
np.random.seed(0)
env = refload.ref_module("DQL").DroneEnv(row_count=G, col_count=G, step_size=1.0, n_anamolous=5, n_drones=D)
agent = ns["Agent"](env.state_size, env.action_size, D)
state = np.reshape(env.reset(), [1, env.state_size])
pos = np.array(env.n_drones_pos).reshape((D, 2))
t0, timed = None, 0
for k in range(T + 40):
    if k == 40:                                                             # memory is past batch_size: every step replays
        t0 = time.perf_counter()
    action = agent.act(state, pos)
    nstate, reward, done = env.step(action)
    reward = reward if not done else 1000
    nstate = np.reshape(nstate, [1, env.state_size])
    npos = np.array(env.n_drones_pos).reshape((D, 2))
    agent.memorize(state, pos, action, reward, nstate, npos, done)
    state, pos = nstate, npos
    if done:
        state = np.reshape(env.reset(), [1, env.state_size])
        pos = np.array(env.n_drones_pos).reshape((D, 2))
        continue
    if len(agent.memory) > 32:
        agent.replay(32)
    timed += k >= 40
dt = time.perf_counter() - t0
print(json.dumps({"impl": "reference DQLearning.py loop (host)", "n_drones": D, "grid": G, "steps": T, "replay_batch": 32,
                  "torch_threads": torch.get_num_threads(), "ms_per_train_step": dt / T * 1e3, "agent_steps_per_s": D * T / dt}))
