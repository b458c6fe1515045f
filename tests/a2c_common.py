"""Shared by tests/golden/a2c/make_a2c_golden.py and tests/test_a2c_*.py: deterministic reference-shaped weights and
tensor digests (so fixtures hold no weight tensors)."""
import os

import numpy as np
import torch

A2C_GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "a2c")


def ref_weights(seed, shapes):
    """state_dict of one reference `Policy`: every tensor ~ U(-b, b), b = 1/sqrt(last dim), drawn key by key (sorted) from
    numpy's legacy MT19937 (frozen stream) — identical on every machine."""
    rs = np.random.RandomState(seed)
    out = {}
    for k in sorted(shapes):
        shape = tuple(shapes[k])
        b = 1.0 / np.sqrt(shape[-1]) if len(shape) > 1 else 0.1
        out[k] = torch.from_numpy(rs.uniform(-b, b, size=shape).astype(np.float32))
    return out


def digest(t):
    """float64 [12]: sum, sum|.|, sum of squares, max|.|, and the first 8 entries of a tensor (zeros for None)."""
    if t is None:
        return np.zeros(12)
    a = t.detach().cpu().double().flatten().numpy()
    head = np.zeros(8)
    head[:min(8, a.size)] = a[:8]
    return np.concatenate([[a.sum(), np.abs(a).sum(), (a * a).sum(), np.abs(a).max()], head])


def policy_shapes(obs_size, n_drones, action_size=4):
    """name -> shape of `Policy(obs_size, n_drones, action_size, 'MLP').state_dict()` (A2C/models.py:33-88)."""
    from swarm_b200.a2c import _mlp_shapes
    sh = {}
    for head, a in (("actor", action_size), ("critic", 1)):
        for k, s in _mlp_shapes(obs_size, n_drones).items():
            sh[f"{head}.{k}"] = s
        sh[f"{head}.dense4.weight"] = (a, 32)
        sh[f"{head}.dense4.bias"] = (a,)
    return sh


def icm_weights(seed, obs_size, action_space=4):
    """{'features' | 'inverse_model' | 'forward_model': state_dict} for the reference ICM's three modules
    (A2C/curiosity.py:57-61), drawn like ref_weights (numpy legacy MT19937, identical everywhere)."""
    from swarm_b200.a2c import BatchedICM
    ref = BatchedICM(obs_size, action_space)
    return {name: ref_weights(seed + 1000 * k, {key: tuple(v.shape) for key, v in getattr(ref, name).state_dict().items()})
            for k, name in enumerate(("features", "inverse_model", "forward_model"))}
