"""N > 1 host logic of the two batched trainers on CPU: world_size-2 gloo processes, each holding half of a minibatch, average
their gradients through `allreduce_grads_` (one all-reduce of the flattened gradient); the result must equal the gradient of the
unsharded minibatch (both losses are means over the batch, shards are equal-sized)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from test_dql_cpu import q_weights, random_batch

R, CC, D, B = 5, 5, 3, 16


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _dql_grads(net, batch, lo, hi):
    from swarm_b200.dql import dql_loss
    t = torch.from_numpy
    net.zero_grad(set_to_none=True)
    dql_loss(net, *(t(a[lo:hi]) for a in batch), 0.95).backward()


def _a2c_case():
    from swarm_b200.a2c import BatchedPolicy
    torch.manual_seed(4)
    pol = BatchedPolicy(R * CC, D, dtype=torch.float64)
    obs = torch.from_numpy(np.random.RandomState(2).choice([0.0, 1.0, 5.0], size=(B, R * CC)))
    return pol, obs


def _a2c_grads(pol, obs, lo, hi):
    pol.zero_grad(set_to_none=True)
    logits, values = pol(obs[lo:hi])
    (torch.log_softmax(logits, -1)[..., 1].mean() + (values ** 2).mean()).backward()   # any batch-mean loss will do
    for prm in pol.dead_parameters():
        if prm.grad is None:
            prm.grad = torch.zeros_like(prm)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from swarm_b200.a2c import A2CTrainer
    from swarm_b200.dql import BatchedDQLAgent, BatchedQNet
    half = B // world
    net = BatchedQNet(R * CC, D, dtype=torch.float64).load_reference(q_weights(9, R * CC, D))
    batch = random_batch(np.random.RandomState(21), B, R, CC, D)
    agent = BatchedDQLAgent.__new__(BatchedDQLAgent)                        # the method only needs .net and .group
    agent.net, agent.group = net, None
    _dql_grads(net, batch, rank * half, (rank + 1) * half)
    agent.allreduce_grads_()
    pol, obs = _a2c_case()
    tr = A2CTrainer.__new__(A2CTrainer)
    tr.policy, tr.group = pol, None
    _a2c_grads(pol, obs, rank * half, (rank + 1) * half)
    tr.allreduce_grads_()
    if rank == 0:
        out.put(({k: p.grad.numpy() for k, p in net.named_parameters()}, {k: p.grad.numpy() for k, p in pol.named_parameters()}))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_rank_gradient_average_equals_unsharded_gradient():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    dql_g, a2c_g = out.get(timeout=150)
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    from swarm_b200.dql import BatchedQNet
    net = BatchedQNet(R * CC, D, dtype=torch.float64).load_reference(q_weights(9, R * CC, D))
    _dql_grads(net, random_batch(np.random.RandomState(21), B, R, CC, D), 0, B)
    for k, p in net.named_parameters():
        np.testing.assert_allclose(dql_g[k], p.grad.numpy(), rtol=1e-10, atol=1e-12, err_msg=k)
    pol, obs = _a2c_case()
    _a2c_grads(pol, obs, 0, B)
    for k, p in pol.named_parameters():
        np.testing.assert_allclose(a2c_g[k], p.grad.numpy(), rtol=1e-10, atol=1e-12, err_msg=k)
