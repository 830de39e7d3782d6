"""Baseline: PyTorch + NCCL all-reduce data-parallel sparse logistic regression.

Same model, data shape, batch and BSP semantics as the parameter-server benchmark, written
the way a PyTorch user would: every rank holds a full replica of the weights, computes its
batch gradient with stock PyTorch ops (gather, segment-sum via index_add_), all-reduces the
dense gradient over NCCL and applies it.  None of this repo's kernels are on this path.
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
        os.environ.setdefault("MASTER_PORT", "29512")
        dist.init_process_group("nccl", rank=rank, world_size=world)
    nnz = 40
    n_local = args.samples // world
    B, K, W = args.batch, args.steps, max(3, args.warmup)
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    # same synthetic shape as minips_b200.models.lr.make_synthetic, built with plain torch
    col = torch.randint(0, args.features, (n_local, nnz), device=dev, generator=g)
    val = torch.randn(n_local, nnz, device=dev, generator=g)
    label = (torch.rand(n_local, device=dev, generator=g) < 0.5).float()
    w = torch.zeros(args.features, device=dev)
    alpha = args.alpha
    cg = torch.Generator().manual_seed(7 + rank)

    def step(start):
        c = col[start:start + B]
        v = val[start:start + B]
        y = label[start:start + B]
        z = (w[c] * v).sum(1)
        pred = torch.sigmoid(z)
        coef = alpha * (y - pred)
        grad = torch.zeros_like(w).index_add_(0, c.reshape(-1), (coef[:, None] * v).reshape(-1))
        if world > 1:
            dist.all_reduce(grad)
        w.add_(grad)
        return torch.nn.functional.binary_cross_entropy_with_logits(z, y, reduction="sum")

    starts = torch.randint(0, max(1, n_local - B), (W + K,), generator=cg).tolist()
    for s in starts[:W]:
        step(s)
    sampler = ClockSampler(local).start()
    sync_barrier(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loss = None
    for s in starts[W:]:
        loss = step(s)
    e1.record()
    sync_barrier(dev)
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    clocks = sampler.stop()
    if rank == 0:
        emit({
            "metric": "sparse_lr_samples_per_sec", "value": world * B * K / (ms / 1e3), "unit": "samples/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "nccl_dp_pytorch",
            "config": {"model": "sparse_logistic_regression", "samples": args.samples, "features": args.features,
                       "nnz_per_row": nnz, "global_batch": B * world, "batch_per_gpu": B, "consistency": "BSP",
                       "parallelism": f"dp{world}_allreduce", "loss_last": float(loss) / B},
            "clocks": clocks, "gpu_launches": 0,
        })
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
