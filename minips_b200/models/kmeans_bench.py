"""`bench.py --app kmeans`: mini-batch k-means samples/s — K=1024, d=128 dense, 100M points
(BASELINE.json config 3), centroid push/pull through the parameter server with the fused
tcgen05 distance GEMM.  One rank per GPU; weak scaling (per-GPU batch fixed)."""
from __future__ import annotations

import torch

K_CLUSTERS = 1024
DIM = 128
N_POINTS = 100_000_000


def run(args, emit):
    import torch.distributed as dist
    from ..parallel.engine import Engine

    engine = Engine.from_env()
    engine.start_everything(num_server_threads_per_node=max(1, 8 // len(engine.nodes)))
    rec = measure(engine, args)
    engine.stop_everything()
    if engine.node.id == 0:
        emit(rec)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def measure(engine, args) -> dict:
    """The k-means benchmark on an already started engine; returns the record (every rank)."""
    from ..parallel.types import MLTask, ModelType, WorkerAlloc
    from ..utils.benchkit import ClockSampler, max_over_ranks, sync_barrier
    from .kmeans import KMeansTrainer, init_random, make_blobs

    world, rank = len(engine.nodes), engine.node.id
    dev = f"cuda:{engine.device}"
    servers = engine.num_server_threads_per_node
    K, D = K_CLUSTERS, DIM
    n_total = args.samples if args.samples != 10_000_000 else N_POINTS
    n_local = n_total // world
    B = args.batch if args.batch != 65536 else 1 << 20
    steps, warm = args.steps, max(3, args.warmup)
    x = make_blobs(n_local, D, K, device=dev, seed=100 + rank, spread=0.2)
    model = ModelType[args.model]
    t_c = engine.create_table(num_dims=K * D, model_type=model, staleness=args.staleness)
    t_n = engine.create_table(num_dims=K, model_type=model, staleness=args.staleness)
    c0 = init_random(x[: 1 << 20], K, seed=1)
    g = torch.Generator().manual_seed(11 + rank)
    res = {}

    def worker(info):
        tc, tn = info.create_kv_client_table(t_c), info.create_kv_client_table(t_n)
        if info.worker_id == 0:
            tc.add_all(c0.reshape(-1).contiguous())  # "set" = add onto zero (kmeans.cpp:185)
        tc.clock()
        tn.clock()
        tr = KMeansTrainer(tc, tn, x, K, alpha=1.0)
        # fused: assign(+pull) / hist / scan / scatter / run-sum / delta+push / clock of both tables, from one host call
        res["launches_per_step"] = 7 if tr._fusable() else 13
        starts = torch.randint(0, max(1, n_local - B), (warm + steps,), generator=g).tolist()
        sse = torch.zeros(1, device=dev)
        for s in starts[:warm]:
            tr.step(s, B, sse)
        sse0 = None
        sampler = ClockSampler(engine.device).start()
        sync_barrier(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i, s in enumerate(starts[warm:]):
            sse.zero_()
            tr.step(s, B, sse)
            if i == 0:
                sse0 = sse.clone()
        e1.record()
        sync_barrier(dev)
        res["ms"] = max_over_ranks(e0.elapsed_time(e1), dev)
        res["sse_first"], res["sse_last"] = float(sse0) / B, float(sse) / B
        if not args.no_e2e:
            # end to end: every step's batch comes from pinned host memory, its SSE goes back
            nb = 8
            host = [x[i * B:(i + 1) * B].cpu().pin_memory() for i in range(nb)]
            stage = [torch.empty(B, D, dtype=torch.bfloat16, device=dev) for _ in range(2)]
            copy_stream = torch.cuda.Stream(dev)
            ready = [torch.cuda.Event() for _ in range(2)]
            free = [None, None]
            host_sse = torch.zeros(steps, dtype=torch.float32).pin_memory()
            dev_sse = torch.zeros(steps, 1, device=dev)
            cur = torch.cuda.current_stream()

            def upload(i):
                k = i % 2
                with torch.cuda.stream(copy_stream):
                    if free[k] is not None:
                        copy_stream.wait_event(free[k])
                    stage[k].copy_(host[i % nb], non_blocking=True)
                    ready[k].record(copy_stream)

            def e2e(nsteps, timed):
                upload(0)
                for i in range(nsteps):
                    if i + 1 < nsteps:
                        upload(i + 1)
                    k = i % 2
                    cur.wait_event(ready[k])
                    tr.x = stage[k]
                    tr.step(0, B, dev_sse[i] if timed else sse)
                    ev = torch.cuda.Event()
                    ev.record(cur)
                    free[k] = ev
                    if timed:
                        host_sse[i:i + 1].copy_(dev_sse[i], non_blocking=True)

            e2e(warm, False)
            sync_barrier(dev)
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            e2e(steps, True)
            t1.record()
            sync_barrier(dev)
            res["e2e_ms"] = max_over_ranks(t0.elapsed_time(t1), dev)
            res["e2e_sse_last"] = float(host_sse[-1]) / B
        res["clocks"] = sampler.stop()

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(n.id, 1) for n in engine.nodes])
    task.set_tables([t_c, t_n])
    task.set_lambda(worker)
    engine.run(task)
    ms = res["ms"]
    rec = {
        "metric": "kmeans_samples_per_sec", "value": world * B * steps / (ms / 1e3), "unit": "samples/s",
        "n_gpus": world, "steps": steps, "warmup": warm, "ms_per_step": ms / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "ours",
        "config": {"model": "minibatch_kmeans", "K": K, "dim": D, "points": n_total, "global_batch": B * world,
                   "batch_per_gpu": B, "consistency": args.model, "server_shards": servers * world,
                   "parallelism": f"dp{world}+ps_range_sharded{servers * world}",
                   "l2_policy": f"inputs larger than L2: every step reads a fresh {B * D * 2 / 1e6:.0f} MB batch of a "
                                f"{n_local * D * 2 / 1e9:.1f} GB per-GPU dataset",
                   "sse_per_point_first": res["sse_first"], "sse_per_point_last": res["sse_last"]},
        "clocks": res["clocks"], "gpu_launches": res["launches_per_step"] * steps,
    }
    if "e2e_ms" in res:
        rec["e2e"] = {"value": world * B * steps / (res["e2e_ms"] / 1e3), "unit": "samples/s",
                      "h2d_bytes_per_step": B * D * 2, "d2h_bytes_per_step": 4, "steps": steps,
                      "ms_per_step": res["e2e_ms"] / steps, "sse_per_point_last": res["e2e_sse_last"]}
    del x
    torch.cuda.empty_cache()
    return rec
