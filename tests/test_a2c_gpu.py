"""GPU: the batched A2C rollout loop (swarm_b200.a2c + swarm_sample_actions / swarm_nstep_returns of libswarm_b200.so) against
the numpy oracle and against traces of the unmodified reference `train()` (tests/golden/a2c/, A2C/train.py:35-157).

Bars: the env trajectory (obs, rewards, dones) and the return scan are bit-exact; quantities that pass through float32
GEMMs / autograd (loss, gradients, updated weights) within 1e-4 relative to the tensor's scale (summation order differs
between the reference's D batch-1 passes and one stacked pass)."""
import os

import numpy as np
import pytest
import torch

import a2c_common as ac
from oracle import a2c_oracle
from swarm_b200 import BatchedDroneEnv
from swarm_b200.a2c import A2CTrainer, BatchedICM, BatchedPolicy, nstep_returns, sample_actions

pytestmark = pytest.mark.gpu
DEV = "cuda"


def load(name):
    z = np.load(os.path.join(ac.A2C_GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def test_nstep_returns_kernel_bit_exact_and_differentiable():
    g = load("rollout_scan")
    rs = np.random.RandomState(3)
    cases = [(g["rewards"], g["dones"], g["last"], float(g["gamma"]), g["returns"])]
    T, D, N = 25, 8, 1000                                                     # ragged N (not a multiple of the block size)
    cases.append((rs.randint(-2, 131, (T, N)).astype(np.float32), (rs.rand(T, N) < 0.1).astype(np.uint8),
                  rs.randn(D, N).astype(np.float32), 0.99, None))
    for rew, done, last, gamma, want in cases:
        ret_o, coef_o = a2c_oracle.nstep_returns_np(rew, done, last, gamma)
        if want is not None:
            assert np.array_equal(ret_o, want)
        lv = torch.from_numpy(last).to(DEV).requires_grad_(True)
        ret = nstep_returns(torch.from_numpy(rew).to(DEV), torch.from_numpy(done).to(DEV), lv, gamma)
        assert np.array_equal(ret.detach().cpu().numpy(), ret_o)              # bit-exact float32
        w = torch.from_numpy(rs.randn(*ret_o.shape).astype(np.float32)).to(DEV)
        (ret * w).sum().backward()
        want_grad = (w.cpu().numpy() * coef_o[:, None, :]).sum(0)
        assert np.allclose(lv.grad.cpu().numpy(), want_grad, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("D,N,A", [(8, 4096, 4), (3, 1001, 4), (5, 257, 6), (1, 33, 2)])
def test_sample_actions_matches_oracle(D, N, A):
    rs = np.random.RandomState(D * 7 + A)
    logits = (rs.randn(D, N, A) * 2).astype(np.float32)
    want, margin = a2c_oracle.sample_actions_np(logits, seed=0x1234567855, step=9, env_id_base=1 << 33)
    got = sample_actions(torch.from_numpy(logits).to(DEV), seed=0x1234567855, step=9, env_id_base=1 << 33).cpu().numpy()
    assert got.shape == (N, D) and got.dtype == np.uint8 and got.max() < A
    bad = got != want
    assert not (bad & (margin > 1e-5)).any(), "device and oracle draws may differ only on a CDF edge (exp rounding)"
    assert bad.mean() < 1e-3
    # a different step or env id base gives a different stream
    other = sample_actions(torch.from_numpy(logits).to(DEV), seed=0x1234567855, step=10, env_id_base=1 << 33).cpu().numpy()
    if N * D >= 1000:
        assert (other != got).mean() > 0.2


def test_sample_actions_distribution():
    p = np.array([0.05, 0.15, 0.3, 0.5], np.float32)
    logits = torch.from_numpy(np.tile(np.log(p), (4, 50000, 1))).to(DEV)
    a = sample_actions(logits, seed=1, step=0).cpu().numpy()
    freq = np.bincount(a.reshape(-1), minlength=4) / a.size
    assert np.abs(freq - p).max() < 5e-3


def _trainer_from_golden(g, grad_norm_limit):
    grid, D, N, T, seed = (int(v) for v in g["meta"])
    env = BatchedDroneEnv(N, row_count=grid, col_count=grid, n_anamolous=5, n_drones=D, step_size=1.0, obs_mode="full",
                          auto_reset=True, seed=1)
    env.set_state(grid=np.stack([g[f"grid0_{n}"] for n in range(N)]), pos=np.stack([g[f"pos0_{n}"] for n in range(N)]).astype(np.float32),
                  t=np.zeros(N, np.int32))
    sh = ac.policy_shapes(grid * grid, D)
    pol = BatchedPolicy(grid * grid, D, 4, device=DEV).load_reference([ac.ref_weights(seed + i, sh) for i in range(D)])
    icm = None
    if "enable_icm" in g and int(g["enable_icm"]):
        icm = BatchedICM(grid * grid, device=DEV).load_reference(ac.icm_weights(seed + 7, grid * grid)).to(DEV)
    tr = A2CTrainer(env, pol, rollout_steps=T, lr=2e-4, gamma=0.99, grad_norm_limit=grad_norm_limit, reset_each_rollout=False, icm=icm)
    tr._obs = env.stage_obs()
    actions = torch.from_numpy(np.stack([g[f"actions_{n}"] for n in range(N)], axis=1)).to(DEV)   # [T, N, D]
    return env, pol, tr, actions, (grid, D, N, T)


def _close(got, want, scale, tol=1e-4):
    return np.abs(np.asarray(got) - np.asarray(want)).max() <= tol * max(scale, 1e-6)


def test_one_env_iteration_equals_reference_train():
    """N = 1: rollout, loss, clipped gradients and the AdamW-updated weights equal one outer iteration of the reference."""
    g = load("train_5x5_d2_n1")
    env, pol, tr, actions, (grid, D, N, T) = _trainer_from_golden(g, float(g["grad_norm_limit"]))
    logits, values, final_values = tr.rollout(actions)
    assert np.array_equal(tr._rew.cpu().numpy()[:, 0], g["rewards_0"].astype(np.float32))
    assert np.array_equal(tr._done.cpu().numpy()[:, 0], g["dones_0"])
    assert np.array_equal(env.stage_obs().cpu().numpy()[0], g["obs_0"][-1].astype(np.float32))
    loss, norms = tr.backward(logits, values, final_values)
    assert abs(float(loss) - float(g["loss_0"])) < 1e-4 * abs(float(g["loss_0"]))
    keys = [str(k) for k in g["keys"]]
    for i in range(D):
        # the reference's clip_grad_norm_ had clip_coef < 1 iff the norm exceeded the limit; the recorded norm is post-clip
        gd = pol.reference_grad_dict(i)
        want = g[f"grad_0_{i}"]
        scale = np.abs(want[:, 3]).max()
        for k, w in zip(keys, want):
            assert _close(ac.digest(gd[k])[[0, 1, 3, 4, 5, 6, 7]], w[[0, 1, 3, 4, 5, 6, 7]], max(scale, abs(w[1]))), (i, k)
    tr.opt.step()
    for i in range(D):
        sd = pol.reference_state_dict(i)
        for k, w in zip(keys, g[f"weight_0_{i}"]):
            # weights move by ~lr per step: compare the first entries at 1e-6 absolute (an lr-sized error would be 2e-4)
            assert np.abs(ac.digest(sd[k])[4:] - w[4:]).max() < 2e-6, (i, k)
            assert abs(ac.digest(sd[k])[0] - w[0]) < 1e-4 * max(1.0, w[1]), (i, k)
    env.close()


def test_batched_gradient_is_mean_of_independent_reference_runs():
    """N = 3 envs from the same weights: the batched (unclipped) gradient is the mean of three reference train() iterations."""
    g = load("train_6x6_d3_n3")
    env, pol, tr, actions, (grid, D, N, T) = _trainer_from_golden(g, float(g["grad_norm_limit"]))
    logits, values, final_values = tr.rollout(actions)
    for n in range(N):
        assert np.array_equal(tr._rew.cpu().numpy()[:, n], g[f"rewards_{n}"].astype(np.float32))
        assert np.array_equal(tr._done.cpu().numpy()[:, n], g[f"dones_{n}"])
    loss, _ = tr.backward(logits, values, final_values)
    want_loss = np.mean([float(g[f"loss_{n}"]) for n in range(N)])
    assert abs(float(loss) - want_loss) < 1e-4 * abs(want_loss)
    keys = [str(k) for k in g["keys"]]
    lin = [0, 4, 5, 6, 7, 8, 9, 10, 11]                                       # digest entries that are linear in the tensor
    for i in range(D):
        want = np.mean([g[f"grad_{n}_{i}"] for n in range(N)], axis=0)
        scale = max(np.abs(g[f"grad_{n}_{i}"][:, 3]).max() for n in range(N))
        gd = pol.reference_grad_dict(i)
        for k, w in zip(keys, want):
            assert _close(ac.digest(gd[k])[lin], w[lin], max(scale, np.abs(w[lin]).max())), (i, k)
    env.close()


def test_sampled_training_runs_and_learns_signal():
    """The full loop with on-device sampling at a cfg4-like shape: finite loss, weights move, launches counted."""
    env = BatchedDroneEnv(512, row_count=10, col_count=10, n_drones=8, step_size=1.0, obs_mode="full", auto_reset=True, seed=3)
    tr = A2CTrainer(env, rollout_steps=5, seed=11)
    before = tr.policy.dense1__weight.detach().clone()
    l0 = env.launch_count
    losses = [float(tr.train_iteration()) for _ in range(3)]
    assert all(np.isfinite(losses))
    assert (tr.policy.dense1__weight.detach() - before).abs().max() > 0
    assert env.launch_count - l0 >= 3 * 5
    assert tr._act.max() < 4 and env.device_error() == 0
    env.close()


def test_icm_bonus_iteration_equals_reference_train():
    """enable_icm: the curiosity bonus (drone 0's logits, s_t fixed at the rollout's first obs) enters the rewards exactly as in
    the reference train() (train.py:85-92); loss and the mean gradient of two independent runs match."""
    g = load("train_icm_5x5_d2_n2")
    env, pol, tr, actions, (grid, D, N, T) = _trainer_from_golden(g, float(g["grad_norm_limit"]))
    logits, values, final_values = tr.rollout(actions)
    for n in range(N):
        bonus = tr._rew.cpu().numpy()[:, n] - g[f"rewards_{n}"].astype(np.float32)      # recorded rewards are the env's own
        assert (bonus > 0).all() and bonus.max() < 50, "the curiosity bonus is a small positive prediction error"
        assert np.array_equal(tr._done.cpu().numpy()[:, n], g[f"dones_{n}"])
    loss, _ = tr.backward(logits, values, final_values)
    want_loss = np.mean([float(g[f"loss_{n}"]) for n in range(N)])
    assert abs(float(loss) - want_loss) < 1e-4 * abs(want_loss)
    keys = [str(k) for k in g["keys"]]
    lin = [0, 4, 5, 6, 7, 8, 9, 10, 11]
    for i in range(D):
        want = np.mean([g[f"grad_{n}_{i}"] for n in range(N)], axis=0)
        scale = max(np.abs(g[f"grad_{n}_{i}"][:, 3]).max() for n in range(N))
        gd = pol.reference_grad_dict(i)
        for k, w in zip(keys, want):
            assert _close(ac.digest(gd[k])[lin], w[lin], max(scale, np.abs(w[lin]).max())), (i, k)
    env.close()


@pytest.mark.parametrize("N,G,H", [(1, 2, 25), (37, 6, 100), (513, 16, 36)])
def test_fused_gru_gates_match_the_torch_expression(N, G, H):
    """swarm_gru_gates_fwd / _bwd against the element-wise definition (torch.nn.GRU cell at h0 = 0, A2C/models.py:104) in
    float64: values within 1e-6, gradients of gi and bias_hh within 1e-5 relative."""
    from swarm_b200.a2c import _GruGates
    gen = torch.Generator(device="cuda").manual_seed(7)
    gi = (torch.randn((N, G * 3 * H), device="cuda", generator=gen) * 1.5).requires_grad_(True)
    bhh = torch.randn((G, 3 * H), device="cuda", generator=gen).requires_grad_(True)
    w = torch.randn((G, N, H), device="cuda", generator=gen)
    h = _GruGates.apply(gi, bhh)
    assert h.shape == (G, N, H) and h.is_contiguous()
    (h * w).sum().backward()
    gi64 = gi.detach().double().requires_grad_(True)
    bh64 = bhh.detach().double().requires_grad_(True)
    g3 = gi64.view(N, G, 3 * H).transpose(0, 1)
    b3 = bh64.unsqueeze(1)
    r = torch.sigmoid(g3[..., :H] + b3[..., :H])
    z = torch.sigmoid(g3[..., H:2 * H] + b3[..., H:2 * H])
    n = torch.tanh(g3[..., 2 * H:] + r * b3[..., 2 * H:])
    h64 = (1.0 - z) * n
    (h64 * w.double()).sum().backward()
    assert torch.allclose(h.double(), h64, rtol=1e-6, atol=1e-6)
    assert torch.allclose(gi.grad.double(), gi64.grad, rtol=1e-5, atol=1e-6)
    assert torch.allclose(bhh.grad.double(), bh64.grad, rtol=1e-5, atol=1e-5 * max(1.0, N ** 0.5))


@pytest.mark.parametrize("shape", [(3, 5, 256), (2, 129, 100), (1, 7, 33)])
def test_fused_relu_l2norm_matches_torch(shape):
    """swarm_relu_l2norm_fwd / _bwd against F.normalize(F.relu(x), dim=-1) in float64 (A2C/models.py:107-108), including a
    row that relu zeroes completely (norm below eps: output 0, gradient blocked by the relu)."""
    import torch.nn.functional as F
    from swarm_b200.a2c import _ReluL2Norm
    gen = torch.Generator(device="cuda").manual_seed(11)
    x = torch.randn(shape, device="cuda", generator=gen)
    x[0, 0] = -x[0, 0].abs() - 0.1                      # an all-negative row
    x = x.requires_grad_(True)
    w = torch.randn(shape, device="cuda", generator=gen)
    o = _ReluL2Norm.apply(x)
    (o * w).sum().backward()
    x64 = x.detach().double().requires_grad_(True)
    o64 = F.normalize(F.relu(x64), dim=-1)
    (o64 * w.double()).sum().backward()
    assert torch.allclose(o.double(), o64, rtol=1e-6, atol=1e-7)
    assert torch.allclose(x.grad.double(), x64.grad, rtol=1e-5, atol=1e-6)
    assert float(o.detach()[0, 0].abs().max()) == 0.0 and float(x.grad[0, 0].abs().max()) == 0.0


def test_fused_policy_equals_float64_policy_at_cfg4_size():
    """BASELINE configs[3] shape (16 384 envs x 8 drones, 10x10): the float32 CUDA policy — fused GRU gates, fused relu +
    normalize, split-K weight gradients — against the same stacked weights evaluated op by op in float64 on the same inputs.
    Outputs within 2e-5.  Gradients of a scalar probe loss: float32 rounding is amplified by the 1 / ||.|| of F.normalize, so
    the yardstick is the op-by-op float32 torch expression (`fused = False`) measured against float64 the same way — the fused
    path may not be further from float64 than 2x that (+1e-4 of the tensor's largest entry)."""
    N, D, H = 16384, 8, 100
    torch.manual_seed(3)
    pol = BatchedPolicy(H, D, 4, device="cuda")
    pol64 = BatchedPolicy(H, D, 4, device="cuda").double()
    with torch.no_grad():
        for a, b in zip(pol64.parameters(), pol.parameters()):
            a.copy_(b.double())
    gen = torch.Generator(device="cuda").manual_seed(5)
    obs = torch.randint(0, 3, (N, H), device="cuda", generator=gen).float()
    obs[obs == 2] = 5.0                                   # cells are 0 / 1 / U
    lw = torch.randn((D, N, 4), device="cuda", generator=gen)
    vw = torch.randn((D, N, 1), device="cuda", generator=gen)

    def grads(p, x, fused):
        p.fused = fused
        for prm in p.parameters():
            prm.grad = None
        lg, vl = p(x)
        ((lg * lw.to(lg.dtype)).sum() + (vl * vw.to(vl.dtype)).sum()).backward()
        return lg.detach(), vl.detach(), [None if prm.grad is None else prm.grad.double().clone() for prm in p.parameters()]

    lg64, vl64, g64 = grads(pol64, obs.double(), False)
    lgu, vlu, gu = grads(pol, obs, False)
    lgf, vlf, gf = grads(pol, obs, True)
    assert torch.allclose(lgf.double(), lg64, rtol=0, atol=2e-5) and torch.allclose(vlf.double(), vl64, rtol=0, atol=2e-5)
    checked = 0
    for (name, _), a64, au, af in zip(pol.named_parameters(), g64, gu, gf):
        if a64 is None or af is None or float(a64.abs().max()) == 0.0:
            continue
        scale = float(a64.abs().max())
        err_f, err_u = float((af - a64).abs().max()) / scale, float((au - a64).abs().max()) / scale
        assert err_f <= 2.0 * err_u + 1e-4, (name, err_f, err_u)
        checked += 1
    assert checked >= 16


@pytest.mark.parametrize("G,N,fin,fout,relu", [(2, 7, 10, 33, True), (4, 1024, 100, 256, True), (3, 65, 32, 32, False),
                                               (8, 2048, 128, 64, 2), (3, 9, 5, 7, 2)])
def test_stacked_linear_epilogue_and_split_k_match_torch(G, N, fin, fout, relu):
    """`_StackedLinear` with the in-place bias (+ relu, or tanh when relu == 2: the deep-Q nets) epilogue (swarm_bias_act, 128-bit and scalar paths) and the split-K
    weight gradient against torch.baddbmm / F.relu autograd on the same float32 inputs."""
    import torch.nn.functional as F
    from swarm_b200.a2c import _StackedLinear
    gen = torch.Generator(device="cuda").manual_seed(21)
    mk = lambda *s: torch.randn(s, device="cuda", generator=gen)
    x, w, b, gw = mk(G, N, fin), mk(G, fout, fin) * 0.1, mk(G, fout), mk(G, N, fout)
    outs = []
    for fused in (True, None):
        xx, ww, bb = (t.clone().requires_grad_(True) for t in (x, w, b))
        if fused:
            y = _StackedLinear.apply(xx, ww, bb, relu, True)
        else:
            y = torch.baddbmm(bb.unsqueeze(1), xx, ww.transpose(1, 2))
            y = torch.tanh(y) if relu == 2 else (F.relu(y) if relu else y)
        (y * gw).sum().backward()
        outs.append((y.detach(), xx.grad, ww.grad, bb.grad))
    for a, r in zip(*outs):
        assert torch.allclose(a, r, rtol=1e-5, atol=1e-5 * max(1.0, float(r.abs().max())))
