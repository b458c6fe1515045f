"""CPU: the batched re-expression of the reference A2C nets / rollout maths (swarm_b200.a2c) against fixtures
frozen from the LIVE reference (tests/golden/a2c/, make_a2c_golden.py) and, when /root/reference is present, against the
reference modules themselves.  Tolerance: float32 GEMM re-association only — 2e-5 absolute on O(0.1) logits / values."""
import os
import sys

import numpy as np
import pytest
import torch

import a2c_common as ac
from oracle import a2c_oracle
from swarm_b200.a2c import BatchedPolicy, a2c_loss

TOL = 2e-5


def load(name):
    z = np.load(os.path.join(ac.A2C_GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def golden_policy(seed, obs_size, n_drones):
    pol = BatchedPolicy(obs_size, n_drones, 4)
    sh = ac.policy_shapes(obs_size, n_drones)
    return pol.load_reference([ac.ref_weights(seed + i, sh) for i in range(n_drones)])


@pytest.mark.parametrize("name", ["policy_5x5_d3", "policy_10x10_d2"])
def test_batched_policy_equals_reference_batch1_calls(name):
    g = load(name)
    grid, D, seed = (int(v) for v in g["meta"])
    pol = golden_policy(seed, grid * grid, D)
    with torch.no_grad():
        logits, values = pol(torch.from_numpy(g["obs"]), torch.from_numpy(g["pos"]))
    assert logits.shape == g["logits"].shape and values.shape == g["values"].shape
    assert np.abs(logits.numpy() - g["logits"]).max() < TOL
    assert np.abs(values.numpy() - g["values"]).max() < TOL


def test_state_dict_round_trip():
    pol = golden_policy(3, 25, 2)
    sh = ac.policy_shapes(25, 2)
    for i in range(2):
        sd, want = pol.reference_state_dict(i), ac.ref_weights(3 + i, sh)
        assert sorted(sd) == sorted(want)
        for k in want:
            assert torch.equal(sd[k], want[k]), k


def test_nstep_oracle_equals_reference_process_rollout():
    g = load("rollout_scan")
    ret, coef = a2c_oracle.nstep_returns_np(g["rewards"], g["dones"], g["last"], float(g["gamma"]))
    assert np.array_equal(ret, g["returns"])                                  # bit-exact: same float32 operation order
    assert np.array_equal(ret - g["values"], g["advantages"])
    # coef is d returns / d last_values; returns is affine in last_values, so a unit step measures it
    eps = 1.0
    ret2, _ = a2c_oracle.nstep_returns_np(g["rewards"], g["dones"], g["last"] + np.float32(eps), float(g["gamma"]))
    assert np.abs((ret2 - ret) / eps - coef[:, None, :]).max() < 1e-3


def test_loss_matches_reference_formula_on_golden_train_case():
    """Replays the recorded obs of a reference train() iteration through the batched nets on CPU and rebuilds the loss with
    the numpy scan: equals the loss the reference logged (train.py:123-135)."""
    g = load("train_5x5_d2_n1")
    grid, D, N, T, seed = (int(v) for v in g["meta"])
    pol = golden_policy(seed, grid * grid, D)
    obs0 = g["grid0_0"].reshape(1, -1).astype(np.float32)
    obs_seq = np.concatenate([obs0, g["obs_0"].astype(np.float32)])          # obs fed at step t (0..T-1) and the final obs
    with torch.no_grad():
        logits, values = pol(torch.from_numpy(obs_seq))                       # rows = time steps here (N = 1 env)
    last = values[:, T:T + 1, 0].numpy()                                      # [D, 1]
    ret, _ = a2c_oracle.nstep_returns_np(g["rewards_0"].astype(np.float32)[:, None], g["dones_0"][:, None], last, 0.99)
    lg = logits[:, :T].permute(1, 0, 2).unsqueeze(2)                          # [T, D, 1, A]
    vl = values[:, :T, 0].t().unsqueeze(2)                                    # [T, D, 1]
    loss = a2c_loss(lg, vl, torch.from_numpy(ret))
    assert abs(float(loss) - float(g["loss_0"])) < 1e-3 * abs(float(g["loss_0"]))


@pytest.mark.skipif(not os.path.exists("/root/reference/A2C/models.py"), reason="live reference only in the build container")
def test_batched_policy_vs_live_reference_modules():
    sys.path.insert(0, "/root/reference/A2C")
    import models as ref_models
    torch.manual_seed(0); np.random.seed(0)
    D, grid, N = 3, 6, 9
    nets = [ref_models.Policy(grid * grid, D, 4, policy_type="MLP") for _ in range(D)]     # the reference's own initialisation
    pol = BatchedPolicy(grid * grid, D, 4).load_reference(nets)
    obs = torch.from_numpy(np.random.choice([0.0, 1.0, 5.0], size=(N, grid * grid)).astype(np.float32))
    pos = torch.rand(N, 2 * D) * grid
    logits, values = pol(obs, pos)
    loss = (logits * torch.linspace(-1, 1, 4)).sum() + values.sum()
    loss.backward()
    ref_loss = 0
    for i, net in enumerate(nets):
        for n in range(N):
            p, v = net(obs[n:n + 1], pos[n:n + 1])
            assert (p[0] - logits[i, n]).abs().max() < TOL and (v[0] - values[i, n]).abs().max() < TOL
            ref_loss = ref_loss + (p * torch.linspace(-1, 1, 4)).sum() + v.sum()
    ref_loss.backward()
    # gradients of the stacked parameters equal the reference nets' accumulated gradients; the dead parameters' are zero there
    for i, net in enumerate(nets):
        sd = dict(net.named_parameters())
        for head, gidx in (("actor", i), ("critic", D + i)):
            for key in ("gru.weight_ih_l0", "attn1.v.weight", "state_input.weight", "dense3.bias", "attn3.v.weight"):
                got = getattr(pol, key.replace(".", "__")).grad[gidx]
                assert (got - sd[f"{head}.{key}"].grad).abs().max() < 1e-4, (head, key)
            for key in ("attn1.q.weight", "attn2.k.bias", "attn3.drone_input.weight", "gru.weight_hh_l0"):
                assert sd[f"{head}.{key}"].grad.abs().max() == 0
        assert (pol.actor_dense4__weight.grad[i] - sd["actor.dense4.weight"].grad).abs().max() < 1e-4
    assert all(p.grad is None for p in pol.dead_parameters())


def test_sampler_oracle_distribution():
    """sample_actions_np draws from softmax(logits): chi-square over 40k draws of a fixed 4-way distribution."""
    logits = np.tile(np.log(np.array([0.1, 0.2, 0.3, 0.4], np.float32)), (4, 10000, 1))
    a, _ = a2c_oracle.sample_actions_np(logits, seed=5, step=3)
    freq = np.bincount(a.reshape(-1), minlength=4) / a.size
    assert np.abs(freq - [0.1, 0.2, 0.3, 0.4]).max() < 0.01


def test_batched_icm_equals_reference_bonus():
    """BatchedICM on N rows == the reference ICM.__call__ at batch 1 per row (curiosity.py:74-86), frozen in icm_6x6.npz."""
    from swarm_b200.a2c import BatchedICM
    g = load("icm_6x6")
    obs_size, seed = (int(v) for v in g["meta"])
    icm = BatchedICM(obs_size).load_reference(ac.icm_weights(seed, obs_size))
    r = icm(torch.from_numpy(g["logits"]), torch.from_numpy(g["s_t"]), torch.from_numpy(g["s_t1"]))
    assert np.abs(r.numpy() - g["bonus"]).max() < 1e-5 * max(1.0, np.abs(g["bonus"]).max())


@pytest.mark.skipif(not os.path.exists("/root/reference/A2C/curiosity.py"), reason="live reference only in the build container")
def test_batched_icm_vs_live_reference():
    sys.path.insert(0, "/root/reference/A2C")
    import curiosity as ref_curiosity
    from swarm_b200.a2c import BatchedICM
    torch.manual_seed(3)
    ref = ref_curiosity.ICM(obs_size=25, action_space=4)                     # the reference's own initialisation
    icm = BatchedICM(25).load_reference(ref)
    s_t, s_t1, lg = torch.rand(6, 25) * 5, torch.rand(6, 25) * 5, torch.randn(6, 4)
    got = icm(lg, s_t, s_t1)
    for n in range(6):
        want = ref(a_t=torch.tensor([n % 4]), a_t_logits=lg[n:n + 1], s_t=s_t[n:n + 1], s_t1=s_t1[n:n + 1])
        assert abs(float(got[n]) - want) < 1e-5 * max(1.0, abs(want))
