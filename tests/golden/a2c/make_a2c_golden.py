"""Freezes outputs of the LIVE, UNMODIFIED reference A2C code (A2C/models.py, A2C/train.py, A2C/noVis_sim.py) into
tests/golden/a2c/*.npz.  Run in the build container only: `python tests/golden/a2c/make_a2c_golden.py`
(/root/reference does not exist on the GPU box).

Weights are never stored: `ref_weights(seed, shape_dict)` (tests/a2c_common.py) regenerates them from numpy's frozen legacy
MT19937 stream, so a fixture holds only inputs and the reference's outputs.

  policy_<shape>.npz   Policy(obs, n_drones, 4, "MLP") batch-1 forward (models.py:44-49,100-117) of D nets on N observation rows
  rollout_scan.npz     process_rollout (train.py:160-183) on synthetic steps with dones -> returns / advantages
  train_<case>.npz     one outer iteration of train() (train.py:35-157) per env: reset states, the sampled actions, rewards,
                       dones, loss, per-parameter gradient digests and post-AdamW weight digests
"""
import logging
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, TESTS)
sys.path.insert(0, os.path.join(os.path.dirname(TESTS), "marl-drones_b200"))
sys.path.insert(0, "/root/reference/A2C")
import a2c_common as ac  # noqa: E402
import models as ref_models  # noqa: E402
import curiosity as ref_curiosity  # noqa: E402
import train as ref_train  # noqa: E402
from noVis_sim import DroneEnv  # noqa: E402


def make_nets(seed, obs_size, n_drones):
    nets = [ref_models.Policy(obs_size, n_drones, 4, policy_type="MLP") for _ in range(n_drones)]
    for i, net in enumerate(nets):
        net.load_state_dict(ac.ref_weights(seed + i, {k: tuple(v.shape) for k, v in net.state_dict().items()}))
    return nets


def policy_case(name, grid, n_drones, n_rows, seed):
    obs_size = grid * grid
    nets = make_nets(seed, obs_size, n_drones)
    rs = np.random.RandomState(seed)
    obs = rs.choice([0.0, 1.0, 5.0], size=(n_rows, obs_size), p=[0.4, 0.5, 0.1]).astype(np.float32)
    pos = rs.randint(0, grid, (n_rows, 2 * n_drones)).astype(np.float32)
    logits = np.zeros((n_drones, n_rows, 4), np.float32)
    values = np.zeros((n_drones, n_rows, 1), np.float32)
    with torch.no_grad():
        for i, net in enumerate(nets):
            for n in range(n_rows):
                p, v = net(torch.from_numpy(obs[n:n + 1]), torch.from_numpy(pos[n:n + 1]))   # batch 1, as train.py:66-67
                logits[i, n], values[i, n] = p[0].numpy(), v[0].numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), obs=obs, pos=pos, logits=logits, values=values,
                        meta=np.array([grid, n_drones, seed]))
    print(name, "logits", logits.std(), "values", values.std())


def scan_case():
    rs = np.random.RandomState(7)
    T, D, N = 12, 3, 9
    gamma = 0.99
    rewards = rs.randint(-1, 131, (T, N)).astype(np.float32)
    dones = (rs.rand(T, N) < 0.2).astype(np.uint8)
    values = rs.randn(T, D, N).astype(np.float32)
    last = rs.randn(D, N).astype(np.float32)
    returns = np.zeros((T, D, N), np.float32)
    adv = np.zeros((T, D, N), np.float32)
    args = types.SimpleNamespace(gamma=gamma)
    for n in range(N):                                                   # one reference rollout per env
        steps = []
        for t in range(T):
            r = torch.tensor([float(rewards[t, n])]).float().unsqueeze(1)                                  # train.py:102
            m = (1.0 - torch.from_numpy(np.array([bool(dones[t, n])], dtype=np.float32))).unsqueeze(1)      # train.py:95-97
            steps.append((r, m, torch.zeros(D, 1), torch.zeros(D, 4), torch.from_numpy(values[t, :, n:n + 1].copy())))
        steps.append((None, None, None, None, torch.from_numpy(last[:, n:n + 1].copy())))
        _, _, _, ret, a = ref_train.process_rollout(args, steps)
        returns[:, :, n] = ret.view(T, D).numpy()
        adv[:, :, n] = a.view(T, D).numpy()
    np.savez_compressed(os.path.join(HERE, "rollout_scan.npz"), rewards=rewards, dones=dones, values=values, last=last,
                        returns=returns, advantages=adv, gamma=np.float64(gamma))
    print("rollout_scan ok")


class RecEnv:
    """The reference env behind a recorder (same object surface train() uses: reset, step, n_drones_pos, action_size)."""

    def __init__(self, **kw):
        self.env = DroneEnv(**kw)
        self.action_size, self.state_size, self.n_drones = self.env.action_size, self.env.state_size, self.env.n_drones
        self.resets, self.actions, self.rewards, self.dones, self.obs = [], [], [], [], []

    @property
    def n_drones_pos(self):
        return self.env.n_drones_pos

    def reset(self):
        o = self.env.reset()
        self.resets.append((np.array(self.env.grid, np.int8).T.copy(), np.array(self.env.n_drones_pos, np.float64)))
        return o

    def step(self, actions):
        o, r, d = self.env.step(actions)
        self.actions.append([int(np.asarray(a).reshape(-1)[0]) for a in actions])
        self.rewards.append(r); self.dones.append(bool(d)); self.obs.append(np.asarray(o, np.int8))
        return o, r, d


class RecOpt:
    """Wraps the reference optimizer: records the (clipped) gradients train() hands to step()."""

    def __init__(self, opt, net):
        self.opt, self.net, self.grads = opt, net, None

    def step(self):
        self.grads = {k: (p.grad.detach().clone() if p.grad is not None else None) for k, p in self.net.named_parameters()}
        self.opt.step()

    def zero_grad(self):
        self.opt.zero_grad()


def det_icm(seed, obs_size):
    """A reference ICM whose three modules carry the deterministic test weights (the reference draws them from torch's RNG)."""
    icm = ref_curiosity.ICM(obs_size=obs_size, action_space=4)
    for name, sd in ac.icm_weights(seed, obs_size).items():
        getattr(icm, name).load_state_dict(sd)
    return icm


def icm_case():
    """ICM.__call__ (curiosity.py:74-86) at batch 1 on N rows: the bonus r_i it returns."""
    rs = np.random.RandomState(5)
    obs_size, N = 36, 8
    icm = det_icm(31, obs_size)
    s_t = rs.choice([0.0, 1.0, 5.0], size=(N, obs_size)).astype(np.float32)
    s_t1 = rs.choice([0.0, 1.0, 5.0], size=(N, obs_size)).astype(np.float32)
    logits = rs.randn(N, 4).astype(np.float32)
    a_t = rs.randint(0, 4, N)
    r = np.zeros(N)
    for n in range(N):
        r[n] = icm(a_t=torch.tensor([a_t[n]], dtype=torch.long), a_t_logits=torch.from_numpy(logits[n:n + 1]),
                   s_t=torch.from_numpy(s_t[n:n + 1]), s_t1=torch.from_numpy(s_t1[n:n + 1]))
    np.savez_compressed(os.path.join(HERE, "icm_6x6.npz"), s_t=s_t, s_t1=s_t1, logits=logits, bonus=r, meta=np.array([obs_size, 31]))
    print("icm_6x6 bonus", r[:4])


def train_case(name, grid, n_drones, n_envs, T, seed, grad_norm_limit, enable_icm=False):
    obs_size = grid * grid
    out = dict(meta=np.array([grid, n_drones, n_envs, T, seed]), grad_norm_limit=np.float64(grad_norm_limit), enable_icm=np.int32(enable_icm))
    keys = None
    tmp = tempfile.mkdtemp()
    cwd = os.getcwd()
    os.chdir(tmp)                                                         # train() writes its log file to the cwd
    try:
        for n in range(n_envs):
            nets = make_nets(seed, obs_size, n_drones)                    # every env starts from the SAME weights
            opts = [RecOpt(torch.optim.AdamW(nets[i].parameters(), lr=2e-4), nets[i]) for i in range(n_drones)]
            np.random.seed(seed * 100 + n)
            torch.manual_seed(seed * 100 + n)
            env = RecEnv(row_count=grid, col_count=grid, step_size=1.0, n_anamolous=5, n_drones=n_drones)
            env.resets.clear()                                            # the constructor's init_env is discarded by train()'s reset
            ref_train.ICM = (lambda obs_size, action_space: det_icm(seed + 7, obs_size))   # same class, deterministic weights
            args = types.SimpleNamespace(enable_icm=enable_icm, policy="MLP", total_steps=T, rollout_steps=T, gamma=0.99, value_coeff=0.5,
                                         entropy_coeff=0.01, grad_acc=1, grad_norm_limit=grad_norm_limit, save_freq=10 ** 9 + 7,
                                         grid_size=grid)
            logging.getLogger().handlers.clear()
            ref_train.train(args, nets, opts, env, obs_size, n_drones)
            assert len(env.resets) == 1 + sum(env.dones), "one reset per outer iteration + one per done"
            assert not any(env.dones), "cases are chosen so that no episode ends inside the rollout"
            out[f"grid0_{n}"], out[f"pos0_{n}"] = env.resets[0]
            out[f"actions_{n}"] = np.array(env.actions, np.uint8)
            out[f"rewards_{n}"] = np.array(env.rewards, np.float64)
            out[f"dones_{n}"] = np.array(env.dones, np.uint8)
            out[f"obs_{n}"] = np.array(env.obs, np.int8)
            loss_line = open(os.path.join(tmp, "A2C_ICM_MLP.log" if enable_icm else "A2C_MLP.log")).read().strip().splitlines()[-1]
            out[f"loss_{n}"] = np.float64(loss_line.split("loss: ")[1].split(",")[0])
            for i in range(n_drones):
                g = opts[i].grads
                keys = keys or sorted(g)
                out[f"grad_{n}_{i}"] = np.stack([ac.digest(g[k]) for k in keys])
                sd = nets[i].state_dict()
                out[f"weight_{n}_{i}"] = np.stack([ac.digest(sd[k]) for k in keys])
                out[f"gradnorm_{n}_{i}"] = np.float64(torch.sqrt(sum((v.double() ** 2).sum() for v in g.values() if v is not None)))
    finally:
        os.chdir(cwd)
    out["keys"] = np.array(keys)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", [float(out[f"loss_{n}"]) for n in range(n_envs)], "gradnorm", [float(out[f"gradnorm_{n}_0"]) for n in range(n_envs)])


if __name__ == "__main__":
    policy_case("policy_5x5_d3", 5, 3, 7, 11)
    policy_case("policy_10x10_d2", 10, 2, 5, 12)
    scan_case()
    train_case("train_5x5_d2_n1", 5, 2, 1, 6, 21, 10.0)          # default clip: post-update weights must match
    train_case("train_6x6_d3_n3", 6, 3, 3, 5, 22, 1e9)           # no clip: batched gradient = mean of 3 reference runs
    icm_case()
    train_case("train_icm_5x5_d2_n2", 5, 2, 2, 5, 23, 1e9, enable_icm=True)   # curiosity bonus in the rewards (drone 0, fixed s_t)
