"""Baseline: PyTorch + NCCL all-reduce data-parallel mini-batch k-means (K=1024, d=128).

Stock PyTorch only: bf16 cuBLAS matmul for the distances, argmin, index_add_ partial sums, one
all-reduce of (sums | counts) per step, replicated centroid update.  None of this repo's kernels.
"""
from __future__ import annotations

import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(args, emit):
    import torch
    import torch.distributed as dist
    from minips_b200.utils.benchkit import ClockSampler, max_over_ranks, sync_barrier

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29513")
        dist.init_process_group("nccl", rank=rank, world_size=world)
    K, D = 1024, 128
    n_total = args.samples if args.samples != 10_000_000 else 100_000_000
    n_local = n_total // world
    B = args.batch if args.batch != 65536 else 1 << 20
    steps, warm = args.steps, max(3, args.warmup)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    x = torch.empty(n_local, D, dtype=torch.bfloat16, device=dev)
    for s in range(0, n_local, 1 << 22):
        m = min(1 << 22, n_local - s)
        x[s:s + m] = torch.randn(m, D, device=dev, generator=g).to(torch.bfloat16)
    centers = x[:K].float().clone()
    counts = torch.zeros(K, device=dev)
    cg = torch.Generator().manual_seed(11 + rank)
    chunk = 1 << 18

    def step(start):
        xb = x[start:start + B]
        cb = centers.to(torch.bfloat16)
        cn = (cb.float() ** 2).sum(1)
        buf = torch.zeros(K, D + 1, device=dev)
        sse = torch.zeros((), device=dev)
        for s in range(0, B, chunk):
            xc = xb[s:s + chunk]
            d = cn[None, :] - 2.0 * (xc @ cb.T).float()
            dmin, a = d.min(1)
            xf = xc.float()
            buf[:, :D].index_add_(0, a, xf)
            buf[:, D].index_add_(0, a, torch.ones_like(dmin))
            sse += ((xf * xf).sum(1) + dmin).clamp_min(0).sum()
        if world > 1:
            dist.all_reduce(buf)
        m = buf[:, D]
        upd = (buf[:, :D] - m[:, None] * centers) / (counts + m).clamp_min(1e-30)[:, None]
        centers.add_(torch.where(m[:, None] > 0, upd, torch.zeros_like(upd)))
        counts.add_(m)
        return sse

    starts = torch.randint(0, max(1, n_local - B), (warm + steps,), generator=cg).tolist()
    for s in starts[:warm]:
        step(s)
    sampler = ClockSampler(local).start()
    sync_barrier(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    sse = None
    for s in starts[warm:]:
        sse = step(s)
    e1.record()
    sync_barrier(dev)
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    clocks = sampler.stop()
    if rank == 0:
        emit({
            "metric": "kmeans_samples_per_sec", "value": world * B * steps / (ms / 1e3), "unit": "samples/s",
            "n_gpus": world, "steps": steps, "warmup": warm, "ms_per_step": ms / steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "nccl_dp_pytorch",
            "config": {"model": "minibatch_kmeans", "K": K, "dim": D, "points": n_total, "global_batch": B * world,
                       "batch_per_gpu": B, "parallelism": f"dp{world}_allreduce",
                       "sse_per_point_last": float(sse) / B},
            "clocks": clocks, "gpu_launches": 0,
        })
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
