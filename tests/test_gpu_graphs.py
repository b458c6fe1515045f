"""GPU: the device entry points under CUDA-graph capture (include/swarm_b200.h: "every device entry point is CUDA-graph
capturable") — step launches of three differently shaped handles (both kernels), a rollout, a masked reset and foreign (torch)
kernels interleaved on ONE captured stream, so that the programmatic-dependent-launch edges between consecutive step kernels and
the plain edges to everything else are all exercised; the graph is replayed with fresh actions and checked against the oracle."""
import numpy as np
import pytest
import torch

from test_gpu_parity import assert_state_equal, make_pair

pytestmark = pytest.mark.gpu

SHAPES3 = [
    dict(row_count=64, col_count=64, n_drones=16, step_size=1.0, obs_mode=1, window=5),          # staged warp kernel, LEAN
    dict(row_count=10, col_count=10, n_drones=8, step_size=1.0),                                 # staged warp kernel, FULL obs
    dict(row_count=24, col_count=20, n_drones=40, step_size=1.0, obs_mode=1, window=5),          # one CTA per env
]


def test_captured_graph_of_mixed_launches_replays_like_the_oracle():
    K, T = 6, 3
    Ns = [96, 50, 21]
    pairs = [make_pair(N, auto_reset=True, seed=40 + j, **kw) for j, (N, kw) in enumerate(zip(Ns, SHAPES3))]
    dev = pairs[0][0].device
    acts = [torch.zeros((K, N, o.D), dtype=torch.uint8, device=dev) for N, (g, o) in zip(Ns, pairs)]
    racts = torch.zeros((T, Ns[0], pairs[0][1].D), dtype=torch.uint8, device=dev)
    obs = [torch.zeros((K,) + tuple(g.obs_shape), dtype=torch.float32, device=dev) for g, _ in pairs]
    rew = [torch.zeros((K, N), dtype=torch.float32, device=dev) for N in Ns]
    done = [torch.zeros((K, N), dtype=torch.uint8, device=dev) for N in Ns]
    robs = torch.zeros((T,) + tuple(pairs[0][0].obs_shape), dtype=torch.float32, device=dev)
    rrew = torch.zeros((T, Ns[0]), dtype=torch.float32, device=dev)
    rdone = torch.zeros((T, Ns[0]), dtype=torch.uint8, device=dev)

    def enqueue():
        for k in range(K):
            for j, (g, _) in enumerate(pairs):            # consecutive step kernels of different handles: dependent launches
                _, r, d, _ = g.step(acts[j][k], out=obs[j][k])
                rew[j][k].copy_(r)                         # torch kernels in between: plain edges
                done[j][k].copy_(d)
        pairs[0][0].rollout(racts, obs_out=robs, reward_out=rrew, done_out=rdone)

    rs = np.random.RandomState(3)

    def fresh_actions():
        hs = [rs.randint(0, 4, tuple(a.shape)).astype(np.uint8) for a in acts]
        hr = rs.randint(0, 4, tuple(racts.shape)).astype(np.uint8)
        for a, h in zip(acts, hs):
            a.copy_(torch.from_numpy(h))
        racts.copy_(torch.from_numpy(hr))
        return hs, hr

    def check(hs, hr, tag):
        torch.cuda.synchronize()
        for j, (g, o) in enumerate(pairs):
            for k in range(K):
                oo, orr, od = o.step(hs[j][k])
                assert np.array_equal(obs[j][k].reshape(Ns[j], -1).cpu().numpy(), oo.astype(np.float32)), (tag, j, k)
                assert np.array_equal(rew[j][k].cpu().numpy(), orr.astype(np.float32)) and np.array_equal(done[j][k].cpu().numpy(), od), (tag, j, k)
        o = pairs[0][1]
        for t in range(T):
            oo, orr, od = o.step(hr[t])
            assert np.array_equal(robs[t].reshape(Ns[0], -1).cpu().numpy(), oo.astype(np.float32)), (tag, "rollout", t)
            assert np.array_equal(rrew[t].cpu().numpy(), orr.astype(np.float32)) and np.array_equal(rdone[t].cpu().numpy(), od), (tag, "rollout", t)
        for g, o in pairs:
            assert_state_equal(g, o, tag)

    # eager once (also warms every lazily set function attribute up before the capture), then capture and replay
    hs, hr = fresh_actions()
    enqueue()
    check(hs, hr, "eager")
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    hs, hr = fresh_actions()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            enqueue()
    torch.cuda.current_stream().wait_stream(side)
    for rep in range(3):                                     # the capture itself does not execute: the first replay runs `hs`
        graph.replay()
        check(hs, hr, f"replay {rep}")
        hs, hr = fresh_actions()
    for g, o in pairs:
        assert np.array_equal(g.stats.cpu().numpy(), o.stats)
        assert g.device_error() == 0
        g.close()
