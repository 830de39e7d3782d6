"""KMeans — mini-batch k-means on the device parameter server (reference apps/kmeans/kmeans.cpp:
154-297 with its flag names, :18-52).  Two tables (centroids K*D, counts K); an init task run by
worker 0 seeds the centres (random | kmeans++ | kmeans_parallel) by adding them onto the zero
table, then every worker iterates: pull, assign (tcgen05 GEMM), partial sums, push, clock.
The sampled-SSE report (kmeans_helper.hpp:236-266) is the exact SSE of the worker's batch here.

    torchrun --nproc-per-node 8 -m minips_b200.apps.kmeans --K=1024 --num_dims=128 --batch_size=1048576
"""
from __future__ import annotations

import sys
import time

import torch


def main(argv=None):
    from ..models.kmeans import INIT_MODES, KMeansTrainer, csr_to_points, make_blobs, pad_points, padded_shape
    from ..ops.sparse import load_libsvm_file
    from ..parallel.engine import Engine
    from ..parallel.types import MLTask, ModelType, WorkerAlloc
    from ..utils.flags import Context, define_kmeans_flags

    ctx = Context.get_instance()
    define_kmeans_flags(ctx)
    ctx.parse(argv)
    engine = Engine.from_env()
    rank, world = engine.node.id, len(engine.nodes)
    dev = f"cuda:{engine.device}"
    K, D = ctx.get_int32("K"), ctx.get_int32("num_dims")
    assert 0 <= ctx.get_int32("kStaleness") <= 5, "kStaleness in [0,5] (kmeans.cpp:59-70)"
    try:
        Kp, Dp = padded_shape(K, D)       # any K, num_dims <= 128: run padded on the tcgen05 tile shapes
    except ValueError as e:
        raise SystemExit(str(e))
    servers = max(1, ctx.get_int32("num_servers_per_node"))
    engine.start_everything(num_server_threads_per_node=servers)
    if ctx.get_string("input"):
        # libsvm samples (kmeans.cpp:154-160), parsed on the device; every node keeps its contiguous share
        csr = load_libsvm_file(ctx.get_string("input"), dev)
        per = csr.n_rows // world
        csr = csr.rows(rank * per, per if rank < world - 1 else csr.n_rows - rank * per)
        x = csr_to_points(csr, D)
        n_local = x.shape[0]
        print(f"[node {rank}] loaded {n_local} points of {D} dims from {ctx.get_string('input')}", flush=True)
    else:
        n_local = ctx.get_int32("num_samples")
        x = pad_points(make_blobs(n_local, D, K, device=dev, seed=ctx.get_int32("seed") + rank))
    model = ModelType[ctx.get_string("kModelType")]
    t_c = engine.create_table(num_dims=K * D, model_type=model, staleness=ctx.get_int32("kStaleness"))
    t_n = engine.create_table(num_dims=K, model_type=model, staleness=ctx.get_int32("kStaleness"))
    engine.barrier()

    # init task: worker 0 only (kmeans.cpp:164-192)
    def init_fn(info):
        tc, tn = info.create_kv_client_table(t_c), info.create_kv_client_table(t_n)
        t0 = time.time()
        c0 = INIT_MODES[ctx.get_string("kmeans_init_mode")](x[:, :D], K, seed=ctx.get_int32("seed"))
        tc.add_all(c0.reshape(-1).contiguous())
        tc.clock()
        tn.clock()
        torch.cuda.synchronize()
        print(f"[init] {ctx.get_string('kmeans_init_mode')} seeding took {(time.time() - t0) * 1e3:.1f} ms")

    init = MLTask()
    init.set_worker_alloc([WorkerAlloc(engine.nodes[0].id, 1)])
    init.set_tables([t_c, t_n])
    init.set_lambda(init_fn)
    engine.run(init)
    engine.barrier()

    B = min(ctx.get_int32("batch_size"), n_local)
    iters = ctx.get_int32("num_iters")
    rep = ctx.get_int32("report_interval")

    def worker(info):
        tc, tn = info.create_kv_client_table(t_c), info.create_kv_client_table(t_n)
        tr = KMeansTrainer(tc, tn, x, K, alpha=ctx.get_double("alpha") if ctx.get_double("alpha") > 0 else 1.0, D=D)
        g = torch.Generator().manual_seed(ctx.get_int32("seed") * 100 + info.worker_id)
        starts = torch.randint(0, max(1, n_local - B + 1), (iters,), generator=g).tolist()
        sse = torch.zeros(1, device=dev)
        t0 = time.time()
        for i, s in enumerate(starts):
            sse.zero_()
            tr.step(s, B, sse)
            if rep > 0 and (i + 1) % rep == 0 and info.worker_id == ctx.get_int32("report_worker"):
                v = float(sse) / B
                ms = (time.time() - t0) * 1e3
                print(f"iter {i + 1}\tSSE/point {v:.5f}\t{ms:.1f} ms")
                if ctx.get_string("report_prefix"):
                    with open(f"{ctx.get_string('report_prefix')}_{rank}", "a") as f:
                        f.write(f"{i + 1}\t{v}\t{ms:.3f}\n")
        torch.cuda.synchronize()
        print(f"[node {rank} worker {info.worker_id}] Total training time: {(time.time() - t0) * 1e3:.1f} ms")

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(n.id, max(1, ctx.get_int32("num_workers_per_node"))) for n in engine.nodes])
    task.set_tables([t_c, t_n])
    task.set_lambda(worker)
    engine.run(task)
    engine.stop_everything()
    return 0


if __name__ == "__main__":
    sys.exit(main())
