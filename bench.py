#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): sparse logistic regression samples/s on the
parameter server — synthetic 10M samples x 1M features (~40 nnz/row), 8 server shards,
BSP — device-timed, max over ranks.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|nccl] [--app lr|kmeans]

N > 1 is launched by torchrun (one rank per GPU).  A "step" is one BSP iteration of every
worker: Get -> forward -> gradient -> Add -> Clock on a batch of --batch rows per GPU
(weak scaling: per-GPU batch fixed, global batch = N * batch; the 10M-row dataset is split
evenly over the N GPUs).  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SAMPLES = 10_000_000
N_FEATURES = 1_000_000
NNZ_PER_ROW = 40
N_SHARDS = 8


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl"])
    ap.add_argument("--app", default="lr", choices=["lr", "kmeans"])
    ap.add_argument("--batch", type=int, default=65536, help="rows per GPU per step")
    ap.add_argument("--samples", type=int, default=N_SAMPLES)
    ap.add_argument("--features", type=int, default=N_FEATURES)
    ap.add_argument("--model", default="BSP", choices=["BSP", "SSP", "ASP"])
    ap.add_argument("--staleness", type=int, default=0)
    ap.add_argument("--alpha", type=float, default=0.01)
    ap.add_argument("--dtype", default="fp64", choices=["fp64", "fp32"],
                    help="value type of the table and of all arithmetic (the reference runs Val = double)")
    ap.add_argument("--repeats", type=int, default=5, help="timed regions of --steps steps; the median is reported")
    ap.add_argument("--no-secondary", action="store_true", help="skip the other-precision LR run and the k-means block")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--timeline", action="store_true", help="also print the in-kernel phase timeline (LR)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = same as --steps")
    return ap.parse_args()


def emit(obj):
    print(json.dumps(obj), flush=True)


def run_reference(args):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    try:
        import ref_run
    except Exception as e:  # noqa: BLE001
        emit({"impl": "reference", "unavailable": f"baseline/ref_run.py not importable: {e}"})
        return
    rec = ref_run.run(args)
    if rec is not None:            # under torchrun only rank 0 reports
        emit(rec)


def base_record(args, world, metric, value, unit, ms_per_step, clocks, config, dtype, extra=None):
    rec = {
        "metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": dtype, "data": "synthetic",
        "impl": args.impl, "config": config, "clocks": clocks,
    }
    if extra:
        rec.update(extra)
    return rec


def measure_lr(engine, args, data, dtype_name, want_e2e, want_timeline=False):
    """One table of `dtype_name`, W warm-up steps, then `--repeats` timed regions of K steps each
    (every region bracketed by barrier + synchronize, CUDA events, max over ranks); the median
    region is the number of record.  Optionally the end-to-end loop through the public API."""
    import statistics
    import torch
    from minips_b200.models.lr import SparseLRTrainer
    from minips_b200.parallel.types import MLTask, ModelType, WorkerAlloc
    from minips_b200.utils.benchkit import ClockSampler, max_over_ranks, sync_barrier
    from minips_b200.utils.loader import AsyncBatchLoader, PackedCsrHost

    world, rank = len(engine.nodes), engine.node.id
    dev = f"cuda:{engine.device}"
    B, K, W, R = args.batch, args.steps, max(3, args.warmup), max(1, args.repeats)
    n_local = data.n_rows
    dtype = torch.float64 if dtype_name == "fp64" else torch.float32
    model = ModelType[args.model]
    tid = engine.create_table(num_dims=args.features, model_type=model, staleness=args.staleness, dtype=dtype)
    g = torch.Generator().manual_seed(7 + rank)
    result = {}

    def starts(n):
        # BatchDataSampler::random_start_point (lib/batch_data_sampler.cpp:42-44) per iteration
        return torch.randint(0, max(1, n_local - B), (n,), generator=g, dtype=torch.int64)

    def worker(info):
        tr = SparseLRTrainer(info.create_kv_client_table(tid), data, args.alpha)
        dense = tr.dense_regime(B)
        result["regime"] = "dense" if dense else "sparse"
        result["protocol"] = (("xchg" if tr.xchg_ok() else "pushpull") if dense else
                              ("bucket" if tr.bucket_ok() else "per-nnz"))

        def run(st_host):
            # dense regime: all iterations inside one persistent kernel; sparse regime (batch touches
            # a small part of a huge table): one fused pull+grad+push kernel + one clock kernel per step
            if dense:
                return tr.train_dense(st_host.to(dev), B)
            out = torch.zeros(len(st_host), 2, device=dev)
            for i, s in enumerate(st_host.tolist()):
                tr.step_sparse(s, B, out[i])
            return out

        sampler = ClockSampler(engine.device).start()         # samples through warm-up and every timed region
        run(starts(W))                                       # warm-up (untimed)
        times, loss = [], None
        for _ in range(R):
            st = starts(K)
            sync_barrier(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            loss = run(st)
            e1.record()
            sync_barrier(dev)
            times.append(max_over_ranks(e0.elapsed_time(e1), dev))
        result["times"] = times
        result["ms"] = statistics.median(times)
        result["loss_last"] = float(loss[-1, 0].item()) / B
        result["launches"] = 1 if dense else (K if result["protocol"] == "bucket" else 2 * K)
        if dense and tr.xchg_ok():
            result["nvls"] = bool(engine.tables[tid].xchg.multicast)
            result["reduce"] = engine.tables[tid].xchg.reduce
            result["bcast"] = engine.tables[tid].xchg.bcast
        if want_e2e and dense:
            Ke = args.e2e_steps or K
            # the data set lives in pinned host memory in the loader's compact form (u8 row lengths, 24-bit
            # feature ids, fp32 values, u8 labels; lossless) when it qualifies, else as plain CSR
            host_csr = data.to("cpu")
            packed = PackedCsrHost.from_csr(host_csr)
            loader = AsyncBatchLoader(packed if packed is not None else host_csr.pin(), B, dev, depth=3,
                                      in_place=tr.xchg_ok())
            result["h2d_format"] = (f"packed: u8 row_len + u{packed.id_bits} col + f32 val + u8 label" +
                                    (" + i64 offset per 8 rows, read in place by the training kernel"
                                     if loader.in_place else ", rebuilt into CSR on the device")
                                    if packed is not None else "csr")
            tr.train_stream(loader, starts(W).tolist())      # warm-up
            etimes, hl = [], None
            for _ in range(R):
                rs = starts(Ke).tolist()
                sync_barrier(dev)
                t0 = torch.cuda.Event(enable_timing=True)
                t1 = torch.cuda.Event(enable_timing=True)
                t0.record()
                hl = tr.train_stream(loader, rs)
                t1.record()
                sync_barrier(dev)
                etimes.append(max_over_ranks(t0.elapsed_time(t1), dev))
            result["e2e_times"] = etimes
            result["e2e_ms"] = statistics.median(etimes)
            result["e2e_steps"] = Ke
            result["h2d"] = loader.bytes_per_batch
            result["e2e_launches_per_step"] = loader.launches_per_batch + 1
            result["e2e_loss_last"] = float(hl[-1, 0]) / B
        result["clocks"] = sampler.stop()
        if want_timeline and not dense and getattr(engine.tables[tid], "bucket", None) is not None and rank == 0:
            t = engine.tables[tid].bucket.timeline.cpu().double()
            names = ["gate", "bucket", "request+wait", "serve+wait", "compute", "delta+wait", "apply"]
            print("timeline us (bucketed step, rank 0, CTA 0, last step): " +
                  ", ".join(f"{n}={v:.1f}" for n, v in zip(names, ((t[1:] - t[:-1]) / 1e3).tolist())),
                  file=sys.stderr)
        if want_timeline and dense:
            # untimed extra run with the in-kernel phase timeline (rank 0 prints the means)
            n_tl = 60
            tl = torch.zeros(n_tl * 8, dtype=torch.int64, device=dev)
            tr.train_dense(starts(n_tl).to(dev), B, timeline=tl)
            torch.cuda.synchronize()
            t = tl.view(n_tl, 8).cpu().double()
            if result["protocol"] == "xchg":
                names = ["broadcast-wait", "zero+rows", "clock+owner-wait", "reduce+broadcast", "signal"]
                d = ((t[:, 1:6] - t[:, 0:5]) / 1e3)[5:].mean(0).tolist()
            else:
                names = ["gate", "pull+sync1", "rows+sync2", "pulled-wait", "push+signal"]
                d = ((t[:, 1:6] - t[:, 0:5]) / 1e3)[5:].mean(0).tolist()
            if rank == 0:
                print(f"timeline us ({dtype_name}, {result['protocol']}, rank 0, CTA 0): " +
                      ", ".join(f"{n}={v:.2f}" for n, v in zip(names, d)) +
                      f", iter={((t[1:, 0] - t[:-1, 0]) / 1e3)[5:].mean().item():.2f}", file=sys.stderr)

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(n.id, 1) for n in engine.nodes])
    task.set_tables([tid])
    task.set_lambda(worker)
    engine.run(task)
    result["value"] = world * B * K / (result["ms"] / 1e3)
    return result


def bench_lr_ours(args):
    import torch.distributed as dist
    from minips_b200.models.lr import make_synthetic
    from minips_b200.parallel.engine import Engine

    engine = Engine.from_env()
    world, rank = len(engine.nodes), engine.node.id
    dev = f"cuda:{engine.device}"
    servers = max(1, N_SHARDS // world)
    engine.start_everything(num_server_threads_per_node=servers)
    n_local = args.samples // world
    B, K = args.batch, args.steps
    data = make_synthetic(n_local, args.features, NNZ_PER_ROW, device=dev, seed=1000 + rank)
    result = measure_lr(engine, args, data, args.dtype, not args.no_e2e, args.timeline)
    secondary = {}
    if not args.no_secondary:
        other = "fp32" if args.dtype == "fp64" else "fp64"
        r2 = measure_lr(engine, args, data, other, False, args.timeline)
        secondary[f"sparse_lr_{other}"] = {
            "metric": "sparse_lr_samples_per_sec", "value": r2["value"], "unit": "samples/s", "dtype": other,
            "ms_per_step": r2["ms"] / K, "repeats_ms": r2["times"], "protocol": r2["protocol"],
            "loss_last": r2["loss_last"], "clocks": r2["clocks"]}
    ms = result["ms"]
    val_note = ("table, replica, accumulators, logits, sigmoid and gradient in " + args.dtype +
                "; sample feature values and labels are stored as fp32 and widened exactly")
    cfg = {
        "model": "sparse_logistic_regression", "samples": args.samples, "features": args.features,
        "nnz_per_row": NNZ_PER_ROW, "server_shards": servers * world, "consistency": args.model,
        "staleness": args.staleness, "global_batch": B * world, "batch_per_gpu": B, "seq_len": None,
        "parallelism": f"dp{world}+ps_range_sharded{servers * world}",
        "precision": val_note,
        "l2_policy": "inputs larger than L2: each step streams a fresh random 21 MB CSR batch of a "
                     f"{n_local * NNZ_PER_ROW * 8 / 1e9:.1f} GB per-GPU dataset; the weight replica is L2-resident by design",
        "regime": result["regime"], "protocol": result["protocol"], "nvls_multicast": result.get("nvls"), "owner_reduce": result.get("reduce"), "owner_broadcast": result.get("bcast"),
        "timed_region": (f"median of {len(result['times'])} regions of K steps, each bracketed by barrier + synchronize; "
                         "one persistent cooperative kernel runs the K iterations of a region (device-side clock gating)"
                         if result["regime"] == "dense" else
                         f"median of {len(result['times'])} regions of K steps; per step " +
                         ("one kernel: keys / values / deltas travel as per-owner messages, Clock inside"
                          if result["protocol"] == "bucket" else
                          "one fused pull+SpMV+gradient+push kernel (per-nnz peer ld / red.add) + one clock kernel")),
        "repeats_ms": result["times"], "loss_last": result["loss_last"],
        "host_numa": engine.numa,
    }
    rec = base_record(args, world, "sparse_lr_samples_per_sec", result["value"], "samples/s", ms / K,
                      result["clocks"], cfg, args.dtype, {"gpu_launches": result["launches"] * len(result["times"])})
    if "e2e_ms" in result:
        Ke = result["e2e_steps"]
        rec["e2e"] = {"value": world * B * Ke / (result["e2e_ms"] / 1e3), "unit": "samples/s",
                      "h2d_bytes_per_step": result["h2d"], "d2h_bytes_per_step": 8,
                      "steps": Ke, "ms_per_step": result["e2e_ms"] / Ke, "repeats_ms": result["e2e_times"],
                      "gpu_launches": Ke * result["e2e_launches_per_step"] * len(result["e2e_times"]),
                      "h2d_format": result.get("h2d_format", "csr"),
                      "loss_last": result["e2e_loss_last"]}
    del data
    import torch
    torch.cuda.empty_cache()
    if not args.no_secondary:
        # BASELINE.json config 3 in the same driver-visible line: mini-batch k-means through the same engine
        from minips_b200.models import kmeans_bench
        try:
            km = kmeans_bench.measure(engine, args.__class__(**{**vars(args), "batch": 65536, "samples": N_SAMPLES,
                                                                "no_e2e": True}))
            secondary["kmeans"] = km
        except Exception as e:  # noqa: BLE001 - the headline must survive a failure of the extra block
            secondary["kmeans"] = {"error": repr(e)}
    if secondary:
        rec["secondary"] = secondary
    engine.stop_everything()
    if rank == 0:
        emit(rec)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "nccl":
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        if args.app == "kmeans":
            import nccl_dp_kmeans
            return nccl_dp_kmeans.run(args, emit)
        import nccl_dp_lr
        return nccl_dp_lr.run(args, emit)
    if args.app == "kmeans":
        from minips_b200.models import kmeans_bench
        return kmeans_bench.run(args, emit)
    return bench_lr_ours(args)


if __name__ == "__main__":
    main()
