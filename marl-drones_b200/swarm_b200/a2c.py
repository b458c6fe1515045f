"""Batched A2C rollout loop over the GPU-resident env (BASELINE config 4).

Re-expression of the reference's `train()` (A2C/train.py:35-157) for N envs at once:

  reference (one env, batch-1 calls)                       here (N envs, one pass)
  ---------------------------------------------------------------------------------------------------------------
  nets[i](obs, drone_pos) for i in drones  train.py:66-73  BatchedPolicy.forward: the 2*D MlpPolicy nets (actor + critic
                                                           per drone, A2C/models.py:33-49,75-117) as stacked weights and
                                                           torch.bmm over [2D, N, .]
  Categorical(softmax(p)).sample()         train.py:68-69  swarm_sample_actions (CUDA, Philox), writes uint8 [N, D]
  env.step(actions); reset on done         train.py:76-80  BatchedDroneEnv.step with auto_reset (one fused kernel)
  process_rollout                          train.py:160-183 swarm_nstep_returns (CUDA reverse scan, float32 order kept)
  loss / backward / clip / AdamW           train.py:123-144 same formulas, mean over envs; per-drone clip_grad_norm_

`MlpPolicy.forward` mixes the rows of a batch (global mean/std, the batch fed to the GRU as a sequence, `query @ key.T`
across the batch — SURVEY §3.4), so it is only meaningful at batch 1, which is how every reference call site uses it.  The
batched form below equals N independent batch-1 calls with the same `state_dict`: row-wise mean / unbiased std, ONE GRU
cell step from h = 0, and `attn_module` reduced to its value projection (softmax over a single score is exactly 1).
The query / key / drone_input parameters are kept (state_dict round trip, AdamW weight decay) but receive zero gradients,
exactly as in the reference.

PyTorch supplies the GEMMs (cuBLAS through torch.bmm) and autograd here; the env step and the two scans are kernels of
libswarm_b200.so.  Tested against the unmodified reference `train()` in tests/test_a2c_*.py.
"""
import ctypes as C
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from ._lib import check

_raw_stream = torch._C._cuda_getCurrentRawStream

# (name in MlpPolicy, in_features, out_features) with H = obs_size, A = action_size / 1   (A2C/models.py:76-88)
_MLP_LINEARS = ("attn1.v", "state_input", "dense1", "attn2.v", "dense2", "dense3", "attn3.v")
_ATTN_DEAD = ("drone_input", "q", "k")     # attn_module members with no effect at batch 1


def _mlp_shapes(H, n_drones):
    """Every parameter of one reference MlpPolicy except dense4: name -> shape."""
    sh = {"gru.weight_ih_l0": (3 * H, H), "gru.weight_hh_l0": (3 * H, H), "gru.bias_ih_l0": (3 * H,), "gru.bias_hh_l0": (3 * H,)}
    for name, fi, fo in (("attn1", H, H), ("attn2", 128, 128), ("attn3", 32, 32)):
        sh[f"{name}.drone_input.weight"] = (fo, 2 * n_drones)
        sh[f"{name}.drone_input.bias"] = (fo,)
        for m in ("q", "k", "v"):
            sh[f"{name}.{m}.weight"] = (fo, fi)
            sh[f"{name}.{m}.bias"] = (fo,)
    for name, fi, fo in (("state_input", H, 256), ("dense1", 256, 128), ("dense2", 128, 64), ("dense3", 64, 32)):
        sh[f"{name}.weight"] = (fo, fi)
        sh[f"{name}.bias"] = (fo,)
    return sh


def _pname(key):
    return key.replace(".", "__")


class BatchedPolicy(nn.Module):
    """The D reference `Policy(obs_size, n_drones, action_size, "MLP")` modules (A2C/models.py:33-49) as ONE module.

    Net index g in [0, 2D): g = i is drone i's actor, g = D + i its critic.  Every MlpPolicy parameter `k` except dense4 is
    stored stacked as `k.replace('.', '__')` with shape [2D, *shape]; dense4 is `actor_dense4__{weight,bias}` [D, A, 32] /
    [D, A] and `critic_dense4__{weight,bias}` [D, 1, 32] / [D, 1].
    """

    def __init__(self, obs_size, n_drones, action_size=4, device=None, dtype=torch.float32):
        super().__init__()
        self.obs_size, self.n_drones, self.action_size = obs_size, n_drones, action_size
        G = 2 * n_drones
        kw = dict(device=device, dtype=dtype)
        self._keys = list(_mlp_shapes(obs_size, n_drones).items())
        shapes = dict(self._keys)
        for key, shape in self._keys:
            # PyTorch's default initialisation of nn.Linear / nn.GRU: U(-1/sqrt(fan), 1/sqrt(fan)), fan = in_features
            # (hidden_size for the GRU); biases use their layer's fan
            wkey = key if len(shape) > 1 else key.rsplit(".", 1)[0] + ".weight"
            fan = obs_size if key.startswith("gru.") else shapes[wkey][-1]
            bound = 1.0 / math.sqrt(fan)
            self.register_parameter(_pname(key), nn.Parameter(torch.empty((G,) + shape, **kw).uniform_(-bound, bound)))
        b4 = 1.0 / math.sqrt(32)
        self.actor_dense4__weight = nn.Parameter(torch.empty((n_drones, action_size, 32), **kw).uniform_(-b4, b4))
        self.actor_dense4__bias = nn.Parameter(torch.empty((n_drones, action_size), **kw).uniform_(-b4, b4))
        self.critic_dense4__weight = nn.Parameter(torch.empty((n_drones, 1, 32), **kw).uniform_(-b4, b4))
        self.critic_dense4__bias = nn.Parameter(torch.empty((n_drones, 1), **kw).uniform_(-b4, b4))
        self.ortho_init_()

    @torch.no_grad()
    def ortho_init_(self):
        """MlpPolicy._init_weights (A2C/models.py:90-98): orthogonal weights (ortho_weights, models.py:9-30: SVD of a
        normal [in, out] matrix, transposed) for state_input, dense1..dense4; drawn per net."""
        def ortho(w):                                   # w [B, out, in]
            a = torch.randn(w.shape[0], w.shape[2], w.shape[1], device=w.device, dtype=torch.float32)
            u, _, vh = torch.linalg.svd(a, full_matrices=False)
            q = u if u.shape == a.shape else vh
            w.copy_(q.transpose(1, 2).to(w.dtype))
        for key in ("state_input", "dense1", "dense2", "dense3"):
            ortho(getattr(self, _pname(key + ".weight")))
        ortho(self.actor_dense4__weight)
        ortho(self.critic_dense4__weight)
        return self

    # ---------------------------------------------------------------------------------- reference state_dict round trip
    @torch.no_grad()
    def load_reference(self, state_dicts):
        """state_dicts: D `Policy.state_dict()`s (keys 'actor.*' / 'critic.*', A2C/models.py:37-38) or Policy modules."""
        assert len(state_dicts) == self.n_drones
        for i, sd in enumerate(state_dicts):
            sd = sd.state_dict() if isinstance(sd, nn.Module) else sd
            for head, g in (("actor", i), ("critic", self.n_drones + i)):
                for key, _ in self._keys:
                    getattr(self, _pname(key))[g].copy_(sd[f"{head}.{key}"])
                getattr(self, f"{head}_dense4__weight")[i].copy_(sd[f"{head}.dense4.weight"])
                getattr(self, f"{head}_dense4__bias")[i].copy_(sd[f"{head}.dense4.bias"])
        return self

    def reference_state_dict(self, i):
        """Drone i's parameters under the reference `Policy.state_dict()` keys (loadable with load_state_dict)."""
        out = {}
        for head, g in (("actor", i), ("critic", self.n_drones + i)):
            for key, _ in self._keys:
                out[f"{head}.{key}"] = getattr(self, _pname(key))[g].detach().clone()
            out[f"{head}.dense4.weight"] = getattr(self, f"{head}_dense4__weight")[i].detach().clone()
            out[f"{head}.dense4.bias"] = getattr(self, f"{head}_dense4__bias")[i].detach().clone()
        return out

    def reference_grad_dict(self, i):
        """Drone i's gradients under the reference `Policy.named_parameters()` keys (None where no gradient exists)."""
        out = {}
        for head, g in (("actor", i), ("critic", self.n_drones + i)):
            for key, _ in self._keys:
                gr = getattr(self, _pname(key)).grad
                out[f"{head}.{key}"] = None if gr is None else gr[g]
            for leaf in ("weight", "bias"):
                gr = getattr(self, f"{head}_dense4__{leaf}").grad
                out[f"{head}.dense4.{leaf}"] = None if gr is None else gr[i]
        return out

    def dead_parameters(self):
        """Parameters that get exact-zero gradients in the reference at batch 1: the attention query / key / drone_input
        (softmax over one score) and the GRU's hidden-side weight (h0 = 0)."""
        dead = [getattr(self, _pname(k)) for k, _ in self._keys if k.split(".")[1] in _ATTN_DEAD and k.startswith("attn")]
        return dead + [self.gru__weight_hh_l0]

    def drone_grad_sqnorm(self):
        """[D] squared gradient norm over everything `nets[i].parameters()` holds (clip_grad_norm_, A2C/train.py:136-139)."""
        D = self.n_drones
        tot = None
        for prm in self.parameters():
            if prm.grad is None:
                continue
            s = prm.grad.float().pow(2).flatten(1).sum(1)
            s = s[:D] + s[D:] if s.numel() == 2 * D else s
            tot = s if tot is None else tot + s
        return tot

    # ---------------------------------------------------------------------------------- forward
    def _lin(self, x, key, relu=False):
        w, b = getattr(self, _pname(key + ".weight")), getattr(self, _pname(key + ".bias"))
        return _StackedLinear.apply(x, w, b, bool(relu), bool(self.fused))

    fused = True   # float32 CUDA tensors go through swarm_gru_gates_* / swarm_relu_l2norm_*; False keeps the op-by-op torch expression (checks)

    def forward(self, obs, drone_pos=None):
        """obs float [N, obs_size] (one row per env) -> (logits [D, N, A], values [D, N, 1]).

        Row n of output i equals `nets[i](obs[n:n+1], drone_pos[n:n+1])` of the reference (A2C/models.py:100-117);
        `drone_pos` only feeds the attention keys, which cannot influence a batch-1 output, and is accepted for signature parity.
        """
        D, H = self.n_drones, self.obs_size
        G = 2 * D
        N = obs.shape[0]
        obs = obs.reshape(N, H).to(self.gru__weight_ih_l0.dtype)
        mu = obs.mean(dim=1, keepdim=True)                                  # models.py:101-103 at batch 1: statistics of the row
        std = obs.std(dim=1, keepdim=True)                                  # unbiased, as torch.Tensor.std()
        x = (obs - mu) / std
        # GRU, one step from h0 = 0 (models.py:104): the hidden-side products vanish, only bias_hh survives
        wih, bih, bhh = self.gru__weight_ih_l0, self.gru__bias_ih_l0, self.gru__bias_hh_l0
        gi = torch.addmm(bih.reshape(1, G * 3 * H), x, wih.reshape(G * 3 * H, H).t())                                     # [N, G*3H]
        if self.fused and gi.is_cuda and gi.dtype == torch.float32:
            h = _GruGates.apply(gi, bhh)                                    # one fused pass each way (swarm_gru_gates_fwd / _bwd)
        else:                                                               # float64 / CPU checks of the same arithmetic
            gi = gi.view(N, G, 3 * H).transpose(0, 1)                       # [G, N, 3H]
            bh = bhh.unsqueeze(1)
            r = torch.sigmoid(gi[..., :H] + bh[..., :H])
            z = torch.sigmoid(gi[..., H:2 * H] + bh[..., H:2 * H])
            n = torch.tanh(gi[..., 2 * H:] + r * bh[..., 2 * H:])
            h = (1.0 - z) * n                                               # + z * h0, h0 = 0
        s = self._lin(h, "attn1.v")                                         # attn_mask == 1  (models.py:67-70)
        s = self._lin(s, "state_input")
        if self.fused and s.is_cuda and s.dtype == torch.float32:
            s = _ReluL2Norm.apply(s)                                        # relu + F.normalize fused (swarm_relu_l2norm_fwd / _bwd)
        else:
            s = F.normalize(F.relu(s), dim=-1)                              # models.py:107-108 (L2 over features, eps 1e-12)
        s = self._lin(s, "dense1", relu=True)
        s = self._lin(s, "attn2.v")
        s = self._lin(s, "dense2", relu=True)
        s = self._lin(s, "dense3", relu=True)
        s = self._lin(s, "attn3.v")
        logits = torch.baddbmm(self.actor_dense4__bias.unsqueeze(1), s[:D], self.actor_dense4__weight.transpose(1, 2))
        values = torch.baddbmm(self.critic_dense4__bias.unsqueeze(1), s[D:], self.critic_dense4__weight.transpose(1, 2))
        return logits, values


# ------------------------------------------------------------------------------------------------ CUDA helpers
def sample_actions(logits, seed, step, env_id_base=0, out=None):
    """softmax + Categorical sample (A2C/train.py:68-69) on the device: logits [D, N, A] float32 -> actions uint8 [N, D]."""
    if not logits.is_cuda:
        raise _lib.SwarmError("sample_actions runs on the GPU only (libswarm_b200.so); there is no CPU fallback")
    D, N, A = logits.shape
    logits = logits.detach()
    if logits.dtype != torch.float32 or not logits.is_contiguous():
        logits = logits.float().contiguous()
    if out is None:
        out = torch.empty((N, D), dtype=torch.uint8, device=logits.device)
    check(_lib.lib().swarm_sample_actions(logits.data_ptr(), D, N, A, int(seed) & (2 ** 64 - 1), int(step) & 0xffffffff,
                                          int(env_id_base), out.data_ptr(), _raw_stream(logits.device.index)))
    return out


class _GruGates(torch.autograd.Function):
    """h = GRU cell gates at h0 = 0 (A2C/models.py:104): gi [N, G*3H] (x W_ih^T + b_ih), bias_hh [G, 3H] -> h [G, N, H]."""

    @staticmethod
    def forward(ctx, gi, bhh):
        G, H3 = bhh.shape
        H, N = H3 // 3, gi.shape[0]
        gi_c, bhh_c = gi.detach().contiguous(), bhh.detach().contiguous()
        h = torch.empty((G, N, H), dtype=torch.float32, device=gi.device)
        check(_lib.lib().swarm_gru_gates_fwd(gi_c.data_ptr(), bhh_c.data_ptr(), N, G, H, h.data_ptr(), _raw_stream(gi.device.index)))
        ctx.save_for_backward(gi_c, bhh_c)
        return h

    @staticmethod
    def backward(ctx, dh):
        gi, bhh = ctx.saved_tensors
        G, H3 = bhh.shape
        H, N = H3 // 3, gi.shape[0]
        dh = dh.contiguous()
        dgi = torch.empty_like(gi)
        dnr = torch.empty((N, G, H), dtype=torch.float32, device=gi.device)
        check(_lib.lib().swarm_gru_gates_bwd(gi.data_ptr(), bhh.data_ptr(), dh.data_ptr(), N, G, H, dgi.data_ptr(), dnr.data_ptr(),
                                             _raw_stream(gi.device.index)))
        dbhh = torch.cat([dgi.view(N, G, 3, H)[:, :, :2].sum(0), dnr.sum(0).unsqueeze(1)], dim=1).reshape(G, H3)
        return dgi, dbhh


class _StackedLinear(torch.autograd.Function):
    """y[g] = x[g] @ w[g].T + b[g] for the 2D stacked nets: x [G, N, in], w [G, out, in], b [G, out] -> y [G, N, out].

    Same forward as `torch.baddbmm`; the backward computes the weight gradient as a split-K batched GEMM.  For N in the tens
    of thousands the plain `dy^T x` is G small (out x in) products with a huge reduction dimension, which cuBLAS runs as a
    handful of 32 x 32 tiles (a few hundred CTAs on 148 SMs); splitting N into S slices gives G * S independent products
    and one cheap sum over S."""

    @staticmethod
    def forward(ctx, x, w, b, relu=False, fused=True):
        """relu: False / 0 none, True / 1 relu, 2 tanh (the deep-Q nets, DQL_model/DQLearning.py:43-48)."""
        G, N, _ = x.shape
        act = int(relu)
        if fused and x.is_cuda and x.dtype == torch.float32:
            y = torch.bmm(x, w.transpose(1, 2))                             # plain GEMM, then bias (+ activation) in one in-place pass
            check(_lib.lib().swarm_bias_act(y.data_ptr(), b.detach().contiguous().data_ptr(), G, N, y.shape[2], act,
                                            _raw_stream(x.device.index)))
        else:
            y = torch.baddbmm(b.unsqueeze(1), x, w.transpose(1, 2))
            if act == 1:
                y = F.relu(y)
            elif act == 2:
                y = torch.tanh(y)
        ctx.act = act
        ctx.save_for_backward(x, w, y if act else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        G, N, fin = x.shape
        fout = w.shape[1]
        dy = dy.contiguous()
        if ctx.act == 1:
            dy = torch.ops.aten.threshold_backward(dy, y, 0)               # relu'(pre) = [y > 0]
        elif ctx.act == 2:
            dy = torch.ops.aten.tanh_backward(dy, y)                        # tanh'(pre) = 1 - y^2
        dx = torch.bmm(dy, w) if ctx.needs_input_grad[0] else None
        S = 1
        while S < 16 and N % (2 * S) == 0 and N // (2 * S) >= 512:
            S *= 2
        if S > 1:
            xc = x.contiguous()
            dw = torch.bmm(dy.view(G * S, N // S, fout).transpose(1, 2), xc.view(G * S, N // S, fin)).view(G, S, fout, fin).sum(1)
        else:
            dw = torch.bmm(dy.transpose(1, 2), x)
        return dx, dw, dy.sum(1), None, None


class _ReluL2Norm(torch.autograd.Function):
    """F.normalize(F.relu(x), dim=-1) (A2C/models.py:107-108), x [..., F] float32 on the GPU."""

    @staticmethod
    def forward(ctx, x):
        xc = x.detach().contiguous()
        out = torch.empty_like(xc)
        Fdim = xc.shape[-1]
        check(_lib.lib().swarm_relu_l2norm_fwd(xc.data_ptr(), xc.numel() // Fdim, Fdim, 1e-12, out.data_ptr(), _raw_stream(x.device.index)))
        ctx.save_for_backward(xc)
        return out

    @staticmethod
    def backward(ctx, g):
        (xc,) = ctx.saved_tensors
        g = g.contiguous()
        dx = torch.empty_like(xc)
        Fdim = xc.shape[-1]
        check(_lib.lib().swarm_relu_l2norm_bwd(xc.data_ptr(), g.data_ptr(), xc.numel() // Fdim, Fdim, 1e-12, dx.data_ptr(), _raw_stream(g.device.index)))
        return dx


class _NStepReturns(torch.autograd.Function):
    """returns[t] = rewards[t] + returns[t+1] * gamma * masks[t]  (A2C/train.py:174), differentiable in last_values."""

    @staticmethod
    def forward(ctx, rewards, dones, last_values, gamma):
        T, N = rewards.shape
        D = last_values.shape[0]
        last = last_values.detach().reshape(D, N).float().contiguous()
        returns = torch.empty((T, D, N), dtype=torch.float32, device=rewards.device)
        coef = torch.empty((T, N), dtype=torch.float32, device=rewards.device)
        check(_lib.lib().swarm_nstep_returns(rewards.data_ptr(), dones.data_ptr(), last.data_ptr(), T, N, D, float(gamma),
                                             returns.data_ptr(), coef.data_ptr(), _raw_stream(rewards.device.index)))
        ctx.save_for_backward(coef)
        ctx.last_shape = last_values.shape
        return returns

    @staticmethod
    def backward(ctx, grad_returns):
        (coef,) = ctx.saved_tensors
        g = (grad_returns * coef.unsqueeze(1)).sum(0)
        return None, None, g.reshape(ctx.last_shape), None


def nstep_returns(rewards, dones, last_values, gamma):
    """rewards float32 [T, N], dones uint8 [T, N], last_values [D, N] (or [D, N, 1]) -> returns [T, D, N]."""
    if not rewards.is_cuda:
        raise _lib.SwarmError("nstep_returns runs on the GPU only (libswarm_b200.so); there is no CPU fallback")
    assert rewards.dtype == torch.float32 and dones.dtype == torch.uint8 and rewards.is_contiguous() and dones.is_contiguous()
    return _NStepReturns.apply(rewards, dones, last_values, gamma)


def a2c_loss(logits, values, returns, value_coeff=0.5, entropy_coeff=0.01):
    """A2C/train.py:121-135 per env, averaged over envs.  logits [T, D, N, A], values / returns [T, D, N].

    The reference's quirks are kept: the policy term multiplies the log-probabilities of ALL actions by the advantage
    (`log_action_probs = log_probs.clone()`, train.py:125; `Variable(advantages)` detaches it there, so the critic learns from
    the value term only), and the entropy term enters with a plus sign.  Every reference mean runs over T*D rows (x A columns) of one env; the extra mean over N envs makes the
    batched gradient the average of N independent reference runs.
    """
    advantages = returns - values                                          # train.py:178
    log_probs = F.log_softmax(logits, dim=-1)
    probs = F.softmax(logits, dim=-1)
    policy_loss = (-log_probs * advantages.detach().unsqueeze(-1)).mean()   # Variable(advantages) detaches (legacy constructor)
    value_loss = advantages.pow(2).mean()
    entropy_loss = (-log_probs * probs).mean()
    return policy_loss + value_loss * value_coeff + entropy_loss * entropy_coeff


class A2CTrainer:
    """Batched `train()` (A2C/train.py:35-157): rollout_steps env steps of N envs, then one loss / backward / clipped AdamW step.

    env: BatchedDroneEnv with obs_mode='full', float32 obs, auto_reset=True (the `if dones: obs = env.reset()` of
    train.py:78-80).  `reset_each_rollout=True` keeps the reference's `obs = env.reset()` at the top of every outer iteration
    (train.py:53).  Optimiser: one fused AdamW over the stacked parameters — element-wise identical to the reference's D
    AdamW instances (main.py:112) — after a per-drone clip_grad_norm_ (train.py:136-139).
    """

    def __init__(self, env, policy=None, rollout_steps=25, lr=2e-4, gamma=0.99, value_coeff=0.5, entropy_coeff=0.01,
                 grad_norm_limit=10.0, seed=0, reset_each_rollout=True, fused_optimizer=True, group=None, icm=None):
        self.env, self.T = env, rollout_steps
        self.N, self.D = env.num_envs, env.n_drones
        self.policy = policy if policy is not None else BatchedPolicy(env.state_size, env.n_drones, env.action_size, device=env.device)
        self.gamma, self.value_coeff, self.entropy_coeff = gamma, value_coeff, entropy_coeff
        self.grad_norm_limit, self.seed, self.reset_each_rollout = grad_norm_limit, seed, reset_each_rollout
        self.opt = torch.optim.AdamW(self.policy.parameters(), lr=lr, fused=bool(fused_optimizer) and env.device.type == "cuda")
        self.icm = icm                  # BatchedICM or None (args.enable_icm, train.py:48-49)
        self.group = group              # torch.distributed group for data-parallel gradient averaging (None + initialised = WORLD)
        self.total_steps = 0            # env steps per env, as train.py:46
        self.sample_step = 0
        dev = env.device
        self._rew = torch.empty((self.T, self.N), dtype=torch.float32, device=dev)
        self._done = torch.empty((self.T, self.N), dtype=torch.uint8, device=dev)
        self._act = torch.empty((self.T, self.N, self.D), dtype=torch.uint8, device=dev)
        self._obs = None

    def rollout(self, actions=None, grad=True):
        """train.py:51-119.  actions: optional uint8 [T, N, D] to replay instead of sampling.  Returns (logits [T, D, N, A],
        values [T, D, N], final_values [D, N]); rewards / dones / actions of the rollout are left in self._rew / _done / _act."""
        env, pol = self.env, self.policy
        if self._obs is None or self.reset_each_rollout:
            self._obs = env.reset()                                         # train.py:53
        obs = self._obs
        curr_state = obs.clone() if self.icm is not None else None      # train.py:57 — never advanced inside the rollout
        logits_l, values_l = [], []
        with torch.set_grad_enabled(grad):
            for t in range(self.T):
                logits, values = pol(obs, env.pos)                          # train.py:66-67, all drones and envs
                if actions is None:
                    sample_actions(logits, self.seed, self.sample_step, out=self._act[t])   # train.py:68-69
                else:
                    self._act[t].copy_(actions[t])
                self.sample_step += 1
                obs, rew, done, _ = env.step(self._act[t])                  # train.py:76-80 (auto-reset returns the reset obs)
                if self.icm is not None:                                    # train.py:85-92: drone 0's logits, s_t = curr_state
                    torch.add(rew, self.icm(logits[0].detach(), curr_state, obs), out=self._rew[t])
                else:
                    self._rew[t].copy_(rew)
                self._done[t].copy_(done)
                logits_l.append(logits)
                values_l.append(values.squeeze(-1))
            _, final_values = pol(obs, env.pos)                             # train.py:109-116
        self._obs = obs
        self.total_steps += self.T
        return torch.stack(logits_l), torch.stack(values_l), final_values.squeeze(-1)

    def backward(self, logits, values, final_values):
        """train.py:119-139: returns, loss, backward, per-drone gradient clipping.  Returns (loss, per-drone gradient norms)."""
        returns = nstep_returns(self._rew, self._done, final_values, self.gamma)
        loss = a2c_loss(logits, values, returns, self.value_coeff, self.entropy_coeff)
        loss.backward()
        for prm in self.policy.dead_parameters():                           # zero (not None) grads: AdamW still decays them
            if prm.grad is None:
                prm.grad = torch.zeros_like(prm)
        self.allreduce_grads_()
        return loss.detach(), self.clip_()

    @torch.no_grad()
    def allreduce_grads_(self):
        """Data-parallel replicas (envs sharded over ranks): average the gradients over ranks — ONE NCCL
        all-reduce of the flattened gradient (~10 MB) per iteration, so every rank applies the mean over all N*world envs."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return
        grads = [p.grad for p in self.policy.parameters() if p.grad is not None]
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, group=self.group)
        flat.div_(dist.get_world_size(self.group))
        off = 0
        for g in grads:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()

    def update(self, logits, values, final_values):
        """train.py:119-144: backward() + the AdamW step of every drone."""
        loss, _ = self.backward(logits, values, final_values)
        self.opt.step()                                                     # train.py:140-141
        self.opt.zero_grad(set_to_none=True)
        return loss

    @torch.no_grad()
    def clip_(self):
        """torch.nn.utils.clip_grad_norm_(nets[i].parameters(), limit) for every drone i (train.py:136-139)."""
        D = self.D
        norm = self.policy.drone_grad_sqnorm().sqrt()
        coef = (self.grad_norm_limit / (norm + 1e-6)).clamp(max=1.0)       # [D]
        coef2 = torch.cat([coef, coef])
        for prm in self.policy.parameters():
            if prm.grad is not None:
                c = coef2 if prm.shape[0] == 2 * D else coef
                prm.grad.mul_(c.view(-1, *([1] * (prm.dim() - 1))).to(prm.grad.dtype))
        return norm

    def train_iteration(self, actions=None):
        logits, values, final_values = self.rollout(actions)
        return self.update(logits, values, final_values)


# ------------------------------------------------------------------------------------------------ ICM
class _IcmFeatures(nn.Module):      # A2C/curiosity.py:7-18
    def __init__(self, obs_size, out_size):
        super().__init__()
        self.dense1, self.dense2, self.dense3 = nn.Linear(obs_size, 128), nn.Linear(128, 64), nn.Linear(64, out_size)

    def forward(self, obs):
        return self.dense3(F.leaky_relu(self.dense2(F.leaky_relu(self.dense1(obs)))))


class _IcmInverse(nn.Module):       # A2C/curiosity.py:21-37 (kept for state_dict parity; it does not influence the bonus)
    def __init__(self, n, action_space):
        super().__init__()
        self.in1, self.in2 = nn.Linear(n, n), nn.Linear(n, n)
        self.dense1, self.dense2, self.dense3 = nn.Linear(2 * n, 32), nn.Linear(32, 16), nn.Linear(16, action_space)

    def forward(self, x1, x2):
        out = torch.cat((F.leaky_relu(self.in1(x1)), F.leaky_relu(self.in2(x2))), dim=-1)
        return self.dense3(F.leaky_relu(self.dense2(F.leaky_relu(self.dense1(out)))))


class _IcmForward(nn.Module):       # A2C/curiosity.py:40-54
    def __init__(self, n, action_space):
        super().__init__()
        self.act_in1, self.act_in2 = nn.Linear(action_space, 16), nn.Linear(16, n)
        self.state_in1, self.dense1 = nn.Linear(n, n), nn.Linear(2 * n, n)

    def forward(self, act, state):
        act = F.leaky_relu(self.act_in2(F.leaky_relu(self.act_in1(act))))
        return self.dense1(torch.cat((F.leaky_relu(self.state_in1(state)), act), dim=-1))


class BatchedICM(nn.Module):
    """The intrinsic-curiosity bonus of A2C/curiosity.py:57-86 for N envs at once, on the obs the env left in HBM.

    Submodule / parameter names equal the reference `ICM`'s (`features`, `inverse_model`, `forward_model`), so
    `load_reference(icm)` copies a reference instance.  As in the reference the module is never optimised (train.py creates
    it without an optimiser; its `.backward()` calls only accumulate unused gradients), so the bonus is a fixed random
    feature-prediction error:  r_i = 10 * 0.5 * mean_k (forward_model(logits_0, phi(s_t)) - phi(s_t1))_k^2   per env
    (every layer is row-wise, so a batch of N rows equals N batch-1 calls).  Reference quirks kept by the trainer: the bonus
    uses drone 0's logits only (train.py:85-92) and `s_t` is the FIRST observation of the rollout (`curr_state` is never
    advanced, train.py:57).
    """

    def __init__(self, obs_size, action_space=4, hidden_size=32, device=None):
        super().__init__()
        self.features = _IcmFeatures(obs_size, hidden_size)
        self.inverse_model = _IcmInverse(hidden_size, action_space)
        self.forward_model = _IcmForward(hidden_size, action_space)
        if device is not None:
            self.to(device)

    @torch.no_grad()
    def load_reference(self, icm):
        """icm: a reference `ICM` instance (plain object holding three nn.Modules) or a dict of their state_dicts."""
        for name in ("features", "inverse_model", "forward_model"):
            sd = icm[name] if isinstance(icm, dict) else getattr(icm, name).state_dict()
            getattr(self, name).load_state_dict(sd)
        return self

    @torch.no_grad()
    def forward(self, a_t_logits, s_t, s_t1):
        """a_t_logits [N, A] (drone 0's policy logits), s_t / s_t1 [N, obs] -> bonus [N] (curiosity.py:74-86)."""
        phi_st, phi_st1 = self.features(s_t), self.features(s_t1)
        pred = self.forward_model(a_t_logits, phi_st)
        return (pred - phi_st1).pow(2).mean(dim=-1) * (0.5 * 10.0)
