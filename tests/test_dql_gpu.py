"""GPU: the batched deep-Q loop (swarm_b200/dql.py) on the CUDA env — memory contents against the oracle's trajectory,
terminal reward, the step cap, greedy action parity with float64, and a replay pass that moves the weights."""
import numpy as np
import pytest
import torch

from oracle.oracle import OracleBatch

pytestmark = pytest.mark.gpu


def _mk(N=64, R=5, Cc=5, D=3, seed=3, **agent_kw):
    from swarm_b200 import BatchedDroneEnv
    from swarm_b200.dql import BatchedDQLAgent
    kw = dict(row_count=R, col_count=Cc, n_drones=D, step_size=1.0, n_anamolous=5)
    env = BatchedDroneEnv(N, reward_mode="DQL", obs_mode="full", auto_reset=True, seed=seed, **kw)
    torch.manual_seed(seed)
    return env, BatchedDQLAgent(env, seed=seed, **agent_kw), kw


def test_memory_holds_the_oracle_trajectory():
    """Every memorised transition equals the oracle stepped with the memorised actions (DQL reward tail,
    DQL_model/noVis_sim.py:138-148; reward 1000 on done, DQLearning.py:171); explore draws stay in {0, 1, 2}."""
    N, T = 64, 160
    env, agent, kw = _mk(N=N, memory_steps=T)
    given = torch.randint(0, 4, (T, N, 3), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    o = OracleBatch(N, reward_mode=1, auto_reset=True, seed=3, scan_mode=1, **kw)
    obs0 = o.reset()                                                         # the agent's `state = env.reset()` (DQLearning.py:162)
    ndone = 0
    for t in range(T):
        own = t % 4 == 0                                                     # act()'s draws never move a drone towards x = 0,
        rew, done = agent.step(None if own else given[t])                    # so most steps replay uniform actions
        a = agent.m_act[t].cpu().numpy()
        assert not own or a.max() <= 2                                       # epsilon = 1: all explore, {0, 1, 2} only
        pos_before = o.pos.copy()
        oo, orr, od = o.step(a)
        assert np.array_equal(agent.m_state[t].cpu().numpy(), obs0)
        assert np.array_equal(agent.m_pos[t].cpu().numpy().astype(np.float64), pos_before)
        assert np.array_equal(agent.m_next[t].cpu().numpy(), oo)
        assert np.array_equal(agent.m_done[t].cpu().numpy(), od.astype(bool))
        assert np.array_equal(rew.cpu().numpy(), np.where(od.astype(bool), 1000.0, orr).astype(np.float32))
        assert np.array_equal(agent.m_npos[t].cpu().numpy().astype(np.float64), o.pos)
        ndone += int(od.sum())
        obs0 = oo.copy()
    assert ndone > 0 and len(agent) == N * T
    env.close()


def test_step_cap_restarts_episodes():
    """`for time in range(2000)` (DQLearning.py:166) with max_steps=7: no env's step counter ever exceeds the cap and
    capped envs start a new episode (episode index advances) although they were not done."""
    env, agent, _ = _mk(N=32, R=10, Cc=10, max_steps=7, memory_steps=8)
    for t in range(30):
        agent.step()
        assert int(env.meta[:, 0].max()) < 7
    assert int(env.meta[:, 3].min()) >= 3                                    # >= 3 restarts each in 30 steps
    assert not bool(agent.m_done.any())                                      # a 10x10 grid is not finished in 7 steps
    env.close()


def test_greedy_actions_match_float64_argmax():
    from swarm_b200.dql import BatchedQNet
    env, agent, _ = _mk(N=256, R=10, Cc=10, D=8)
    for _ in range(5):
        agent.step()
    net64 = BatchedQNet(env.state_size, env.n_drones, dtype=torch.float64, device=env.device)
    net64.load_state_dict({k: v.double() for k, v in agent.net.state_dict().items()})
    q32 = agent.net(agent._cur, env.pos)
    q64 = net64(agent._cur, env.pos)
    torch.testing.assert_close(q32.double(), q64, rtol=1e-4, atol=1e-5)
    a = agent.act(agent._cur, env.pos, infer=True).clone()
    top2 = q64.topk(2, dim=-1).values
    clear = ((top2[..., 0] - top2[..., 1]) > 1e-4).t()                       # ignore float32 near-ties
    assert torch.equal(a[clear].long(), q64.argmax(-1).t()[clear])
    env.close()


def test_train_step_replays_and_learns():
    """After the memory exceeds batch_size every train_step runs replay(): finite loss, weights move, epsilon decays
    (DQLearning.py:135-136,186-187); with gamma = 0 (fixed targets) the loss of a fixed
    minibatch falls when it is replayed repeatedly — the bootstrapped loss has moving targets and need not."""
    env, agent, _ = _mk(N=128, memory_steps=16, batch_size=512)
    w0 = agent.net.dense3__weight.detach().clone()
    losses = []
    for t in range(12):
        _, _, loss = agent.train_step()
        assert (loss is None) == (t < 4)                                     # len(memory) > batch_size from the 5th step on
        if loss is not None:
            losses.append(float(loss))
    assert np.isfinite(losses).all() and not torch.equal(w0, agent.net.dense3__weight)
    assert abs(agent.epsilon - 0.995 ** len(losses)) < 1e-12
    from swarm_b200.dql import dql_loss
    batch = agent.sample(512)
    fixed = lambda: dql_loss(agent.net, *batch, 0.0)                         # noqa: E731  gamma = 0: targets do not move
    first = float(fixed().detach())
    opt = torch.optim.AdamW(agent.net.parameters(), lr=1e-3)
    for _ in range(60):
        opt.zero_grad()
        l = fixed()
        l.backward()
        opt.step()
    assert float(l.detach()) < first
    env.close()
