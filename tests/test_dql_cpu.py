"""CPU: the batched deep-Q nets and replay arithmetic (swarm_b200/dql.py) against the LIVE unmodified reference
DQL_model/DQLearning.py where it is present, and against a batch-1 restatement everywhere (travels to the GPU box)."""
import random

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import refload
from a2c_common import ref_weights
from swarm_b200.dql import BatchedQNet, _q_shapes, dql_loss, dql_targets

live = pytest.mark.skipif(not refload.have_reference(), reason="live reference not present")


def q_weights(seed, S, D, A=4):
    sh = {}
    for name, fi, fo in _q_shapes(S, D, A):
        sh[f"{name}.weight"], sh[f"{name}.bias"] = (fo, fi), (fo,)
    return [ref_weights(seed + 17 * i, sh) for i in range(D)]


def batch1_q(sd, state, pos):
    """DQLearning.py:36-50 restated op by op for ONE sample (state [1, S], pos [D, 2]) -> [1, A]."""
    lin = lambda x, n: F.linear(x, sd[f"{n}.weight"].to(x.dtype), sd[f"{n}.bias"].to(x.dtype))   # noqa: E731
    s = (state - state.mean()) / state.std()
    out = torch.cat([torch.tanh(lin(s, "state_input")), torch.tanh(lin(pos.flatten().unsqueeze(0), "drone_input"))], dim=-1)
    out = F.normalize(out)
    return lin(torch.tanh(lin(torch.tanh(lin(out, "dense1")), "dense2")), "dense3")


def random_batch(rs, B, R, Cc, D):
    state = rs.choice([0.0, 1.0, 5.0], size=(B, R * Cc), p=[0.3, 0.6, 0.1]).astype(np.float32)
    nstate = np.where(rs.rand(B, R * Cc) < 0.05, 0.0, state).astype(np.float32)
    pos = np.stack([rs.randint(0, R, (B, D)), rs.randint(0, Cc, (B, D))], -1).astype(np.float32)
    npos = np.clip(pos + rs.randint(-1, 2, pos.shape), 0, min(R, Cc) - 1).astype(np.float32)
    act = rs.randint(0, 3, (B, D)).astype(np.uint8)
    rew = rs.choice([-1.0, 0.0, 1.0, 11.0, 31.0, 1000.0], size=B).astype(np.float32)
    done = rew == 1000.0
    return state, pos, act, rew, nstate, npos, done


@pytest.mark.parametrize("R,Cc,D", [(5, 5, 3), (10, 10, 8), (4, 7, 2)])
def test_qnet_equals_batch1_restatement(R, Cc, D):
    rs = np.random.RandomState(5)
    sds = q_weights(11, R * Cc, D)
    net = BatchedQNet(R * Cc, D, dtype=torch.float64).load_reference(sds)
    state, pos, *_ = random_batch(rs, 9, R, Cc, D)
    q = net(torch.from_numpy(state), torch.from_numpy(pos))
    assert q.shape == (D, 9, 4)
    for i in range(D):
        for n in range(9):
            want = batch1_q(sds[i], torch.from_numpy(state[n:n + 1]).double(), torch.from_numpy(pos[n]).double())
            torch.testing.assert_close(q[i, n:n + 1], want, rtol=1e-12, atol=1e-12)
    for i in range(D):                                                       # state_dict round trip (Agent.save / load)
        back = net.reference_state_dict(i)
        assert all(torch.equal(back[k].float(), sds[i][k]) for k in sds[i])


@live
@pytest.mark.parametrize("R,Cc,D", [(5, 5, 3), (6, 4, 5)])
def test_qnet_vs_live_reference_agent_model(R, Cc, D):
    """float32, same weights: BatchedQNet rows == the reference's AgentModel.forward on every sample (1e-6 relative)."""
    ns = refload.ref_dql_namespace()
    rs = np.random.RandomState(8)
    sds = q_weights(3, R * Cc, D)
    net = BatchedQNet(R * Cc, D).load_reference(sds)
    state, pos, *_ = random_batch(rs, 7, R, Cc, D)
    q = net(torch.from_numpy(state), torch.from_numpy(pos)).detach()
    for i in range(D):
        ref = ns["AgentModel"](R * Cc, D, 4)
        ref.load_state_dict(sds[i])
        for n in range(7):
            want = ref(state[n:n + 1], pos[n]).detach()
            torch.testing.assert_close(q[i, n:n + 1], want, rtol=1e-5, atol=1e-6)


@live
def test_replay_gradient_is_mean_of_reference_fit_gradients():
    """dql_loss's gradient for net i == the mean over the minibatch of the gradients the reference's replay() hands to
    `_fit` (targets built by the reference's own lines incl. `target_f[0][action] = target` with the action LIST,
    DQLearning.py:123-133), all evaluated at the same weights."""
    ns = refload.ref_dql_namespace()
    R, Cc, D, B = 5, 5, 3, 6
    rs = np.random.RandomState(21)
    sds = q_weights(9, R * Cc, D)
    agent = ns["Agent"](R * Cc, 4, D)
    for i in range(D):
        agent.models[i].load_state_dict(sds[i])
    state, pos, act, rew, nstate, npos, done = random_batch(rs, B, R, Cc, D)
    grads = [{k: torch.zeros_like(p) for k, p in agent.models[i].named_parameters()} for i in range(D)]
    for b in range(B):
        action = [int(a) for a in act[b]]
        for i in range(D):
            model = agent.models[i]
            target = float(rew[b])
            if not done[b]:
                target = rew[b] + agent.gamma * np.amax(model(nstate[b:b + 1], npos[b])[0].detach().numpy())
            target_f = model(state[b:b + 1], pos[b]).detach().numpy()
            target_f[0][action] = target
            model.zero_grad()                                               # _fit without the optimizer step (:81-86)
            loss = agent.mse_loss(model(state[b:b + 1], pos[b]), torch.tensor(target_f, dtype=torch.float))
            loss.backward()
            for k, p in model.named_parameters():
                grads[i][k] += p.grad / B
    net = BatchedQNet(R * Cc, D).load_reference(sds)
    t = torch.from_numpy
    dql_loss(net, t(state), t(pos), t(act), t(rew), t(nstate), t(npos), t(done), agent.gamma).backward()
    for i in range(D):
        for k, g in grads[i].items():
            got = getattr(net, k.replace(".", "__")).grad[i]
            torch.testing.assert_close(got, g, rtol=2e-4, atol=1e-6 * max(1.0, float(g.abs().max())))


def test_targets_write_every_listed_action():
    q = torch.arange(2 * 2 * 4, dtype=torch.float32).reshape(2, 2, 4)
    qn = torch.ones(2, 2, 4) * torch.tensor([1.0, 2.0, 3.0, 4.0])
    tf = dql_targets(q, qn, torch.tensor([[0, 2], [1, 1]]), torch.tensor([10.0, 1000.0]), torch.tensor([False, True]), 0.5)
    assert tf[0, 0].tolist() == [12.0, 1.0, 12.0, 3.0] and tf[1, 0].tolist() == [12.0, 9.0, 12.0, 11.0]
    assert tf[0, 1].tolist() == [4.0, 1000.0, 6.0, 7.0] and tf[1, 1].tolist() == [12.0, 1000.0, 14.0, 15.0]


@live
def test_reference_explore_branch_never_draws_last_action():
    """Pins the quirk the batched act() keeps: `random.randrange(0, action_size - 1)` (DQLearning.py:105) is uniform on
    {0, 1, 2} for 4 actions."""
    ns = refload.ref_dql_namespace()
    agent = ns["Agent"](25, 4, 3)
    random.seed(0)
    np.random.seed(0)
    seen = {int(a) for _ in range(300) for a in agent.act(np.ones((1, 25), np.float32), np.zeros((3, 2)))}
    assert seen == {0, 1, 2}
