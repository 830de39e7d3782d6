#!/usr/bin/env python
"""Aggregate host->device ceiling with N ranks copying at the same time (the end-to-end LR path is bound
by it): every rank streams 256 MiB pinned buffers to its GPU for ~1 s, all ranks start together;
prints the per-rank and the aggregate rate.  Launch like the bench:
    python -m torch.distributed.run --nproc-per-node N tools/h2d_ceiling.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl")
    from minips_b200.utils.numa import bind_to_gpu_node
    numa = bind_to_gpu_node(dev)
    n = 256 << 20
    host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(2)]
    for h in host:
        h.fill_(1)
    d = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(2)]
    for i in range(4):
        d[i % 2].copy_(host[i % 2], non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 40
    e0.record()
    for i in range(iters):
        d[i % 2].copy_(host[i % 2], non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    gbs = n * iters / (e0.elapsed_time(e1) * 1e-3) / 1e9
    t = torch.tensor([gbs], device="cuda", dtype=torch.float64)
    lo, total = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(total, op=dist.ReduceOp.SUM)
    if rank == 0:
        print(f"h2d ceiling, {world} rank(s) at once: slowest rank {float(lo):.1f} GB/s, "
              f"aggregate {float(total):.1f} GB/s (numa binding: {numa})")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
