"""CPU restatement of the A2C rollout helpers — TEST INFRASTRUCTURE (only tests/, smoke() and bench.py's CPU legs import it).

  nstep_returns_np   process_rollout's reverse scan (reference A2C/train.py:160-183) in float32, the reference's operation
                     order.  Pinned: tests/test_a2c_cpu.py checks it against tests/golden/a2c/rollout_scan.npz, frozen from the
                     live reference process_rollout.
  sample_actions_np  the spec of swarm_sample_actions: inverse-CDF draw from softmax(logits) with Philox4x32-10 uniforms.
                     The reference samples with torch's global generator (Categorical(probs).sample(), train.py:68-69), whose
                     stream cannot be reproduced on the device: the SAMPLER's parity is unpinned (distribution checked
                     statistically); everything downstream of the sampled actions is pinned by replaying recorded actions.
"""
import ctypes as C

import numpy as np

from . import oracle


def nstep_returns_np(rewards, dones, last_values, gamma):
    """rewards f32 [T, N], dones u8 [T, N], last_values f32 [D, N] -> (returns f32 [T, D, N], coef f32 [T, N])."""
    T, N = rewards.shape
    g = np.float32(gamma)
    ret = last_values.astype(np.float32).copy()
    coef = np.ones(N, np.float32)
    returns = np.zeros((T,) + ret.shape, np.float32)
    coefs = np.zeros((T, N), np.float32)
    for t in reversed(range(T)):
        mask = (np.float32(1.0) - dones[t].astype(np.float32))
        ret = rewards[t][None, :] + (ret * g) * mask[None, :]       # train.py:174, float32 at every step
        coef = (coef * g) * mask
        returns[t], coefs[t] = ret, coef
    return returns, coefs


def philox_uniforms(seed, step, env_ids, n_drones):
    """u [N, D] in [0, 1): word (d % 4) of Philox4x32-10(counter = (env lo, env hi, step, d // 4), key = seed)."""
    L = oracle.lib()
    fn = L.oracle_philox4x32_10
    fn.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    fn.restype = None
    key = (C.c_uint32 * 2)(seed & 0xffffffff, (seed >> 32) & 0xffffffff)
    out = (C.c_uint32 * 4)()
    u = np.zeros((len(env_ids), n_drones), np.float32)
    for k, e in enumerate(env_ids):
        e = int(e) & (2 ** 64 - 1)
        for blk in range((n_drones + 3) // 4):
            ctr = (C.c_uint32 * 4)(e & 0xffffffff, e >> 32, step & 0xffffffff, blk)
            fn(ctr, key, out)
            for j in range(4):
                if blk * 4 + j < n_drones:
                    u[k, blk * 4 + j] = np.float32(out[j] >> 8) * np.float32(1.0 / 16777216.0)
    return u


def sample_actions_np(logits, seed, step, env_id_base=0):
    """logits f32 [D, N, A] -> (actions u8 [N, D], margin f32 [N, D]); margin = distance of u*sum to the nearest CDF edge
    (a device result may differ only where the margin is within exp() rounding)."""
    D, N, A = logits.shape
    u = philox_uniforms(seed, step, [env_id_base + n for n in range(N)], D)
    l = logits.astype(np.float32)
    e = np.exp(l - l.max(-1, keepdims=True)).astype(np.float32)
    cum = np.zeros_like(e)
    acc = np.zeros(e.shape[:-1], np.float32)
    for k in range(A):
        acc = (acc + e[..., k]).astype(np.float32)
        cum[..., k] = acc
    thr = (u.T * cum[..., -1]).astype(np.float32)                    # [D, N]
    a = (cum[..., :-1] <= thr[..., None]).sum(-1)
    margin = np.abs(cum[..., :-1] - thr[..., None]).min(-1) if A > 1 else np.ones_like(thr)
    return a.T.astype(np.uint8), margin.T
