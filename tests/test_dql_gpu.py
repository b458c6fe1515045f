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


@pytest.mark.parametrize("N,D,fa,fb", [(1000, 8, 64, 64), (37, 3, 64, 64), (5, 2, 8, 4), (64, 4, 128, 128)])
def test_fused_embedding_matches_torch(N, D, fa, fb):
    """swarm_embed_tanh_l2norm_fwd / _bwd against F.normalize(cat(tanh, tanh)) autograd on the same float32 inputs, with
    float64 as the yardstick (DQLearning.py:43-46); includes an all-zero row (norm below eps)."""
    import torch.nn.functional as F
    from swarm_b200.dql import _EmbedTanhL2Norm
    gen = torch.Generator(device="cuda").manual_seed(3)
    a = torch.randn((N, D, fa), device="cuda", generator=gen) * 2
    b = torch.randn((N, D, fb), device="cuda", generator=gen) * 2
    a[0, 0], b[0, 0] = 0.0, 0.0
    gw = torch.randn((D, N, fa + fb), device="cuda", generator=gen)
    ba, bb_ = torch.randn((D, fa), device="cuda", generator=gen), torch.randn((D, fb), device="cuda", generator=gen)
    ba[0], bb_[0] = 0.0, 0.0                                                 # keeps row (0, 0) at zero
    ref = lambda x, y, p, q: F.normalize(torch.cat([x + p, y + q], dim=-1).transpose(0, 1).tanh(), dim=-1)   # noqa: E731
    res = []
    for fn, dt in ((_EmbedTanhL2Norm.apply, torch.float32), (ref, torch.float32), (ref, torch.float64)):
        aa, bb, pp, qq = (t.to(dt).clone().requires_grad_(True) for t in (a, b, ba, bb_))
        o = fn(aa, bb, pp, qq)
        (o * gw.to(dt)).sum().backward()
        res.append((o.detach().double(), aa.grad.double(), bb.grad.double(), pp.grad.double(), qq.grad.double()))
    for got, t32, t64 in zip(*res):
        scale = max(1.0, float(t64.abs().max()))
        err_f, err_t = float((got - t64).abs().max()) / scale, float((t32 - t64).abs().max()) / scale
        assert err_f <= 2.0 * err_t + 2e-6, (err_f, err_t)
    assert torch.equal(res[0][0][0, 0], torch.zeros(fa + fb, dtype=torch.float64, device="cuda"))


def test_fused_qnet_equals_unfused():
    from swarm_b200.dql import BatchedQNet, dql_loss
    torch.manual_seed(0)
    net = BatchedQNet(100, 8, device="cuda")
    gen = torch.Generator(device="cuda").manual_seed(5)
    B = 4096
    state = torch.randint(0, 2, (B, 100), device="cuda", generator=gen).float()
    nstate = state * (torch.rand((B, 100), device="cuda", generator=gen) > 0.05)
    pos = torch.randint(0, 10, (B, 8, 2), device="cuda", generator=gen).float()
    act = torch.randint(0, 3, (B, 8), device="cuda", generator=gen, dtype=torch.uint8)
    rew = torch.randint(-1, 32, (B,), device="cuda", generator=gen).float()
    done = torch.rand(B, device="cuda", generator=gen) < 0.05
    outs = []
    for fused in (True, False):
        net.fused = fused
        net.zero_grad(set_to_none=True)
        q = net(state, pos)
        loss = dql_loss(net, state, pos, act, rew, nstate, pos, done, 0.95)
        loss.backward()
        outs.append([q.detach(), loss.detach()] + [p.grad.clone() for p in net.parameters()])
    for a, r in zip(*outs):
        torch.testing.assert_close(a, r, rtol=2e-4, atol=2e-5 * max(1.0, float(r.abs().max())))


@pytest.mark.parametrize("rows,F", [(1000, 100), (7, 25), (33, 4096), (5, 2)])
def test_row_standardize_matches_torch(rows, F):
    from swarm_b200.dql import row_standardize
    gen = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randint(0, 6, (rows, F), device="cuda", generator=gen).float()
    x[0] = torch.arange(F, device="cuda").float()
    got = row_standardize(x).double()
    x64 = x.double()
    want64 = (x64 - x64.mean(1, keepdim=True)) / x64.std(1, keepdim=True)
    want32 = ((x - x.mean(1, keepdim=True)) / x.std(1, keepdim=True)).double()
    ok = torch.isfinite(want64)
    assert torch.equal(torch.isfinite(got), ok)                              # constant rows: 0 / 0 like torch
    err_f, err_t = float((got - want64)[ok].abs().max()), float((want32 - want64)[ok].abs().max())
    assert err_f <= 2.0 * err_t + 1e-6, (err_f, err_t)
