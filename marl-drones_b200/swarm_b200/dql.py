"""Batched deep-Q loop over the GPU-resident env — the DQL counterpart of a2c.py (the second caller the north star names).

Re-expression of `DQL_model/DQLearning.py` for N envs at once:

  reference (one env, batch-1 calls)                          here (N envs, one pass)
  ------------------------------------------------------------------------------------------------------------------
  AgentModel x n_drones            DQLearning.py:27-50,70-79   BatchedQNet: the D nets as stacked weights, torch.bmm over
                                                               [D, N, .]; row-wise mean / unbiased std (batch 1 in the
                                                               reference, so its global statistics are the row's)
  Agent.act                        DQLearning.py:96-110        BatchedDQLAgent.act: ONE uniform draw per env decides
                                                               explore / exploit for all its drones; exploring drones draw
                                                               from {0 .. action_size-2} (`randrange(0, action_size - 1)`
                                                               never yields the last action — kept); else first argmax
  env.step; reward = 1000 if done  DQLearning.py:169-171       BatchedDroneEnv.step (reward_mode 'DQL', auto_reset: the
                                                               reference breaks and resets on done, :184-185) + the
                                                               caller-side terminal reward; `for time in range(2000)` =
                                                               masked reset of envs whose step counter reaches max_steps
  deque(maxlen=3000) / memorize    DQLearning.py:58,89-94      device ring of `memory_steps` step-slots x N transitions
  replay: random.sample, per-sample,
  per-drone target + _fit          DQLearning.py:112-136,81-87 one minibatch pass: target = r (+ gamma max_a Q_i(s')) unless
                                                               done, written at EVERY index in the sample's action list
                                                               (`target_f[0][action] = target` with `action` the list of
                                                               all drones' actions — kept), MSE mean over actions, mean
                                                               over the minibatch, ONE AdamW step, epsilon decay

The reference applies `batch_size * n_drones` AdamW steps one after another inside `replay`; the batched loop takes one step
along the mean of those per-sample gradients evaluated at the same weights (tests/test_dql_cpu.py checks that identity
against the live reference).  PyTorch supplies the small GEMMs and autograd; the env step is the CUDA kernel of
libswarm_b200.so.  The nets are CPU-constructible so the arithmetic can be pinned without a GPU; the agent needs the env
and therefore a GPU.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

class _EmbedTanhL2Norm(torch.autograd.Function):
    """F.normalize(cat(tanh(a + bias_a), tanh(b + bias_b)), dim=-1) for the stacked nets (DQLearning.py:43-46) in one pass each
    way: a [N, D, fa], b [N, D, fb] float32 CUDA (the shared input-layer products WITHOUT bias — a broadcast bias inside addmm
    costs a full-size fill and a read of it), bias_a [D, fa], bias_b [D, fb] -> [D, N, fa + fb]
    (swarm_embed_tanh_l2norm_fwd / _bwd)."""

    @staticmethod
    def forward(ctx, a, b, bias_a, bias_b):
        from . import _lib
        a, b, bias_a, bias_b = a.contiguous(), b.contiguous(), bias_a.detach().contiguous(), bias_b.detach().contiguous()
        N, D, fa = a.shape
        fb = b.shape[2]
        out = torch.empty((D, N, fa + fb), dtype=torch.float32, device=a.device)
        _lib.check(_lib.lib().swarm_embed_tanh_l2norm_fwd(a.data_ptr(), b.data_ptr(), bias_a.data_ptr(), bias_b.data_ptr(), N, D, fa, fb,
                                                          1e-12, out.data_ptr(), torch._C._cuda_getCurrentRawStream(a.device.index)))
        ctx.save_for_backward(a, b, bias_a, bias_b)
        return out

    @staticmethod
    def backward(ctx, g):
        from . import _lib
        a, b, bias_a, bias_b = ctx.saved_tensors
        N, D, fa = a.shape
        fb = b.shape[2]
        g = g.contiguous()
        da, db = torch.empty_like(a), torch.empty_like(b)
        _lib.check(_lib.lib().swarm_embed_tanh_l2norm_bwd(a.data_ptr(), b.data_ptr(), bias_a.data_ptr(), bias_b.data_ptr(), g.data_ptr(),
                                                          N, D, fa, fb, 1e-12, da.data_ptr(), db.data_ptr(),
                                                          torch._C._cuda_getCurrentRawStream(a.device.index)))
        return da, db, da.sum(0), db.sum(0)


def row_standardize(x):
    """(x - mean) / std per row, unbiased std — `(state - state.mean()) / state.std()` at batch 1 (DQLearning.py:40-42); one
    kernel (swarm_row_standardize) for float32 CUDA rows, the torch expression otherwise.  No gradient (inputs are data)."""
    if x.is_cuda and x.dtype == torch.float32 and not x.requires_grad:
        from . import _lib
        x = x.contiguous()
        out = torch.empty_like(x)
        _lib.check(_lib.lib().swarm_row_standardize(x.data_ptr(), x.shape[0], x.shape[1], out.data_ptr(),
                                                    torch._C._cuda_getCurrentRawStream(x.device.index)))
        return out
    return (x - x.mean(dim=1, keepdim=True)) / x.std(dim=1, keepdim=True)


# (name in AgentModel, in_features, out_features); S = state_size, D = n_drones, A = action_size   (DQLearning.py:30-34)
def _q_shapes(S, D, A):
    return (("state_input", S, 64), ("drone_input", 2 * D, 64), ("dense1", 128, 64), ("dense2", 64, 32), ("dense3", 32, A))


class BatchedQNet(nn.Module):
    """The D `AgentModel` nets of one reference `Agent` with stacked parameters `<layer>__weight [D, out, in]`,
    `<layer>__bias [D, out]`.  `load_reference` / `reference_state_dict` convert from / to the reference's per-drone
    `state_dict` files (`Agent.save` / `load`, DQLearning.py:138-144)."""

    def __init__(self, state_size, n_drones, action_size=4, device=None, dtype=torch.float32):
        super().__init__()
        self.state_size, self.n_drones, self.action_size = state_size, n_drones, action_size
        for name, fi, fo in _q_shapes(state_size, n_drones, action_size):
            bound = 1.0 / math.sqrt(fi)                                     # nn.Linear's default init range
            w = torch.empty((n_drones, fo, fi), device=device, dtype=dtype).uniform_(-bound, bound)
            b = torch.empty((n_drones, fo), device=device, dtype=dtype).uniform_(-bound, bound)
            self.register_parameter(f"{name}__weight", nn.Parameter(w))
            self.register_parameter(f"{name}__bias", nn.Parameter(b))

    @torch.no_grad()
    def load_reference(self, state_dicts):
        assert len(state_dicts) == self.n_drones
        for name, _, _ in _q_shapes(self.state_size, self.n_drones, self.action_size):
            for kind in ("weight", "bias"):
                prm = getattr(self, f"{name}__{kind}")
                prm.copy_(torch.stack([sd[f"{name}.{kind}"].to(prm.device, prm.dtype) for sd in state_dicts]))
        return self

    def reference_state_dict(self, i):
        return {f"{name}.{kind}": getattr(self, f"{name}__{kind}")[i].detach().clone()
                for name, _, _ in _q_shapes(self.state_size, self.n_drones, self.action_size) for kind in ("weight", "bias")}

    fused = True   # float32 CUDA tensors: plain batched GEMM + swarm_bias_act (bias + tanh in place) and split-K weight gradients

    def _lin(self, x, name, act=0):
        from .a2c import _StackedLinear
        return _StackedLinear.apply(x, getattr(self, f"{name}__weight"), getattr(self, f"{name}__bias"), act, bool(self.fused))

    def _lin_shared(self, x, name, bias=True):
        """-> [N, D, out]: the layer of every net applied to the same rows."""
        w, b = getattr(self, f"{name}__weight"), getattr(self, f"{name}__bias")
        D, fo, fi = w.shape
        if not bias:
            return torch.mm(x, w.reshape(D * fo, fi).t()).view(x.shape[0], D, fo)
        return torch.addmm(b.reshape(1, D * fo), x, w.reshape(D * fo, fi).t()).view(x.shape[0], D, fo)

    def forward(self, state, drone_pos):
        """state [N, S] (any real dtype), drone_pos [N, D, 2] -> Q [D, N, A]; row n of net i equals
        `models[i](state[n:n+1], drone_pos[n])` of the reference (DQLearning.py:36-50)."""
        D = self.n_drones
        dt = self.state_input__weight.dtype
        N = state.shape[0]
        s = state.reshape(N, self.state_size).to(dt)
        fused = self.fused and s.is_cuda and dt == torch.float32
        s = row_standardize(s) if fused else (s - s.mean(dim=1, keepdim=True)) / s.std(dim=1, keepdim=True)   # :40-42 at batch 1
        p = drone_pos.reshape(N, 2 * D).to(dt)
        # the two input layers see the same rows in every net: ONE [N, in] x [in, D*64] product each instead of D
        if fused:                                                           # bias + tanh + cat + F.normalize + [N, D] -> [D, N] in one pass
            out = _EmbedTanhL2Norm.apply(self._lin_shared(s, "state_input", False), self._lin_shared(p, "drone_input", False),
                                         self.state_input__bias, self.drone_input__bias)
        else:
            ea, eb = self._lin_shared(s, "state_input"), self._lin_shared(p, "drone_input")
            out = F.normalize(torch.cat([ea, eb], dim=-1).transpose(0, 1).tanh(), dim=-1)   # :43-46 (L2 over features, eps 1e-12)
        out = self._lin(out, "dense1", 2)                                   # :47-48 torch.tanh(dense(.))
        out = self._lin(out, "dense2", 2)
        return self._lin(out, "dense3")


def dql_targets(q_state, q_next, actions, rewards, dones, gamma):
    """`replay`'s target tensor for every sample and net (DQLearning.py:123-133).

    q_state, q_next [D, B, A] (detached Q of the sample's state / next state under net i), actions int [B, D] (the memorised
    action LIST of the step), rewards [B], dones bool [B] -> target_f [D, B, A]: Q_i(state) with `target` written at every
    index that appears in the action list."""
    D, B, A = q_state.shape
    target = rewards.to(q_state.dtype).unsqueeze(0) + gamma * q_next.max(dim=-1).values * (~dones.bool()).to(q_state.dtype).unsqueeze(0)
    hit = torch.zeros((B, A), dtype=torch.bool, device=q_state.device)
    hit.scatter_(1, actions.long(), True)
    return torch.where(hit.unsqueeze(0), target.unsqueeze(-1), q_state)


def dql_loss(net, state, pos, actions, rewards, next_state, next_pos, dones, gamma):
    """Sum over the D nets of the minibatch mean of the reference's per-sample `MSELoss(preds, target_f)` (:81-87): its
    gradient for net i is the mean over the minibatch of the gradients `_fit` would apply one by one."""
    preds = net(state, pos)
    with torch.no_grad():                                                   # :124-131: Q(s') and the detached Q(s) the targets start from
        target_f = dql_targets(preds.detach(), net(next_state, next_pos), actions, rewards, dones, gamma)
    return ((preds - target_f) ** 2).mean(dim=(1, 2)).sum()


class BatchedDQLAgent:
    """`Agent` + the episode loop of DQLearning.py:147-190 on a BatchedDroneEnv (reward_mode='DQL', obs_mode='full', float32
    obs, auto_reset=True).  Hyper-parameters default to the reference's (:59-63); `memory_steps` step-slots of N transitions
    replace `deque(maxlen=3000)`; `batch_size` transitions are drawn without replacement (`random.sample`)."""

    def __init__(self, env, net=None, gamma=0.95, epsilon=1.0, epsilon_min=0.01, epsilon_decay=0.995, lr=1e-4,
                 memory_steps=64, batch_size=None, max_steps=2000, terminal_reward=1000.0, seed=0, group=None):
        self.env, self.N, self.D = env, env.num_envs, env.n_drones
        dev = env.device
        self.net = net if net is not None else BatchedQNet(env.state_size, env.n_drones, env.action_size, device=dev)
        self.gamma, self.epsilon, self.epsilon_min, self.epsilon_decay = gamma, epsilon, epsilon_min, epsilon_decay
        self.opt = torch.optim.AdamW(self.net.parameters(), lr=lr, fused=dev.type == "cuda")          # :76-79
        self.K = memory_steps
        self.batch_size = batch_size if batch_size is not None else 32 * self.N                       # 32 per env (:14)
        self.max_steps, self.terminal_reward = max_steps, terminal_reward
        self.group = group              # torch.distributed group of data-parallel replicas (envs sharded over ranks)
        self.gen = torch.Generator(device=dev)
        self.gen.manual_seed(seed)
        S, N, D, K = env.state_size, self.N, self.D, self.K
        self.m_state = torch.empty((K, N, S), dtype=torch.float32, device=dev)
        self.m_next = torch.empty((K, N, S), dtype=torch.float32, device=dev)
        self.m_pos = torch.empty((K, N, D, 2), dtype=torch.float32, device=dev)
        self.m_npos = torch.empty((K, N, D, 2), dtype=torch.float32, device=dev)
        self.m_act = torch.empty((K, N, D), dtype=torch.uint8, device=dev)
        self.m_rew = torch.empty((K, N), dtype=torch.float32, device=dev)
        self.m_done = torch.empty((K, N), dtype=torch.bool, device=dev)
        self.steps = 0                                                      # env steps taken (slots filled = min(steps, K))
        self._cur = torch.empty((N, S), dtype=torch.float32, device=dev)
        self._act = torch.empty((N, D), dtype=torch.uint8, device=dev)
        env.reset(out=self._cur)                                            # :162

    def __len__(self):
        return min(self.steps, self.K) * self.N

    @torch.no_grad()
    def act(self, state, pos, infer=False, out=None):
        """:96-110.  Returns uint8 [N, D]."""
        out = self._act if out is None else out
        greedy = self.net(state, pos).argmax(dim=-1).t().to(torch.uint8)   # first maximal index, as np.argmax
        if infer:
            return out.copy_(greedy)
        explore = torch.rand(self.N, device=state.device, generator=self.gen) <= self.epsilon
        rnd = torch.randint(0, self.env.action_size - 1, (self.N, self.D), device=state.device, generator=self.gen, dtype=torch.uint8)
        return out.copy_(torch.where(explore.unsqueeze(1), rnd, greedy))

    @torch.no_grad()
    def step(self, actions=None):
        """One pass of the inner loop body (:167-178) for all envs: act, env.step, terminal reward, memorise; then the
        2000-step cap as a masked reset.  actions: optional integer [N, D] to take instead of act()'s.  Returns
        (rewards, dones) of the step (views of the ring slot)."""
        env, s = self.env, self.steps % self.K
        self.m_state[s].copy_(self._cur)
        self.m_pos[s].copy_(env.pos)
        if actions is None:
            self.act(self._cur, env.pos, out=self.m_act[s])
        else:
            self.m_act[s].copy_(actions)
        _, rew, done, _ = env.step(self.m_act[s], out=self.m_next[s])
        d = done.bool()
        self.m_done[s].copy_(d)
        torch.where(d, torch.full_like(rew, self.terminal_reward), rew, out=self.m_rew[s])             # :171
        self.m_npos[s].copy_(env.pos)
        capped = (env.meta[:, 0] >= self.max_steps) & ~d                    # auto_reset already restarted the done envs
        env.reset(mask=capped, out=self._cur)                               # emits the obs of every env: also the state copy
        self.steps += 1
        return self.m_rew[s], self.m_done[s]

    def sample(self, batch_size=None):
        """`random.sample(self.memory, batch_size)` (:113): indices into the filled slots without replacement."""
        B = min(batch_size or self.batch_size, len(self))
        idx = torch.randperm(len(self), device=self.m_rew.device, generator=self.gen)[:B]
        f = lambda a: a.reshape(-1, *a.shape[2:])[idx]                      # noqa: E731
        return tuple(f(a) for a in (self.m_state, self.m_pos, self.m_act, self.m_rew, self.m_next, self.m_npos, self.m_done))

    def replay(self, batch_size=None):
        """:112-136 as one minibatch pass + one AdamW step; epsilon decays once per call (:135-136)."""
        state, pos, act, rew, nstate, npos, done = self.sample(batch_size)
        loss = dql_loss(self.net, state, pos, act, rew, nstate, npos, done, self.gamma)
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        self.allreduce_grads_()
        self.opt.step()
        if self.epsilon > self.epsilon_min:
            self.epsilon *= self.epsilon_decay
        return loss.detach()

    @torch.no_grad()
    def allreduce_grads_(self):
        """Data-parallel replicas: ONE all-reduce of the flattened gradient per replay, so every rank steps along the mean
        over all ranks' minibatches (same scheme as A2CTrainer.allreduce_grads_)."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return
        grads = [p.grad for p in self.net.parameters() if p.grad is not None]
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, group=self.group)
        flat.div_(dist.get_world_size(self.group))
        off = 0
        for g in grads:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()

    def train_step(self):
        """One iteration of the reference's inner loop: step, then `if len(memory) > batch_size: replay` (:186-187)."""
        rew, done = self.step()
        loss = self.replay() if len(self) > self.batch_size else None
        return rew, done, loss
