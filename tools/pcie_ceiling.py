#!/usr/bin/env python
"""The box's host-copy ceiling, the roofline of bench.py's `e2e` number: pinned D2H (and H2D) copies of an obs-sized buffer,
one rank at a time and all ranks at once, no kernel running.  Launch like the bench:

    python tools/pcie_ceiling.py                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 tools/pcie_ceiling.py

Rank 0 prints one JSON object: per-rank GB/s alone, per-rank GB/s with every rank copying, the aggregate, CPU affinity."""
import json
import os
import time

import torch
import torch.distributed as dist


def timed_copies(dst, src, n, dev):
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(n):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize(dev)
    return (time.perf_counter() - t0) / n


def main():
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    out = {"world": world, "cpu_affinity": sorted(os.sched_getaffinity(0))[:4] + ["...", len(os.sched_getaffinity(0))]}
    for name, nbytes in (("obs_float32_105MB", 65536 * 16 * 25 * 4), ("obs_uint8_26MB", 65536 * 16 * 25)):
        d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        for _ in range(3):
            h.copy_(d, non_blocking=True); d.copy_(h, non_blocking=True)
        torch.cuda.synchronize(dev)
        alone_d2h = alone_h2d = 0.0
        for r in range(world):                       # one rank at a time
            if world > 1:
                dist.barrier()
            if r == rank:
                alone_d2h = nbytes / timed_copies(h, d, 10, dev) / 1e9
                alone_h2d = nbytes / timed_copies(d, h, 10, dev) / 1e9
        if world > 1:
            dist.barrier()
        both_d2h = nbytes / timed_copies(h, d, 10, dev) / 1e9     # every rank at once
        if world > 1:
            dist.barrier()
        both_h2d = nbytes / timed_copies(d, h, 10, dev) / 1e9
        vals = torch.tensor([alone_d2h, alone_h2d, both_d2h, both_h2d], dtype=torch.float64, device=dev)
        allv = [torch.zeros_like(vals) for _ in range(world)] if world > 1 else [vals]
        if world > 1:
            dist.all_gather(allv, vals)
        m = torch.stack(allv).cpu()
        out[name] = {"bytes": nbytes,
                     "d2h_alone_gbs_per_rank": [round(float(x), 1) for x in m[:, 0]],
                     "h2d_alone_gbs_per_rank": [round(float(x), 1) for x in m[:, 1]],
                     "d2h_concurrent_gbs_per_rank": [round(float(x), 1) for x in m[:, 2]],
                     "h2d_concurrent_gbs_per_rank": [round(float(x), 1) for x in m[:, 3]],
                     "d2h_concurrent_aggregate_gbs": round(float(m[:, 2].sum()), 1),
                     "h2d_concurrent_aggregate_gbs": round(float(m[:, 3].sum()), 1)}
        del d, h
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
