"""LRExample — sparse logistic regression by mini-batch SGD on the device parameter server.

CLI-compatible with the reference executable (apps/logistic_regression/lr_example.cpp:20-56 flag
names; launch one process per GPU with torchrun or scripts/logistic_regression.py):

    torchrun --nproc-per-node 8 -m minips_b200.apps.lr_example --kModelType=BSP --num_dims=1000000 \
        --batch_size=65536 --num_iters=1000 --alpha=0.1

Structure mirrors Training() (lr_example.cpp:97-385): load (or synthesise) this node's samples,
create the table, then per worker: iterate Get -> gradient -> Add -> Clock (fused kernels),
checkpoint every `checkpoint_interval` iterations, optional report file, straggler injection,
final accuracy on the local data and the total training time.
"""
from __future__ import annotations

import os
import random
import sys
import time

import torch


def _train(engine, ctx, job=None):
    """One generation of Training() (lr_example.cpp:97-385) on a started-up process group: create the
    table, restore it when resuming, run the workers, tear the engine down."""
    from ..models.lr import SparseLRTrainer, make_synthetic
    from ..ops.sparse import load_libsvm_file
    from ..parallel.types import MLTask, ModelType, StorageType, WorkerAlloc

    engine.context.update(ctx.values)
    dev = f"cuda:{engine.device}"
    servers = max(1, ctx.get_int32("num_servers_per_node"))
    engine.start_everything(num_server_threads_per_node=servers)
    rank, world = engine.rank_index, len(engine.nodes)
    resume = ctx.get_bool("use_weight_file") or (job is not None and job.resume)

    num_dims = ctx.get_int32("num_dims")
    if ctx.get_string("input"):
        data = load_libsvm_file(ctx.get_string("input"), dev)
        # every node parses the file and keeps its contiguous share of the rows
        per = data.n_rows // world
        data = data.rows(rank * per, per if rank < world - 1 else data.n_rows - rank * per)
        data.col = data.col.contiguous()
        data.val = data.val.contiguous()
        bad = int(((data.col < 0) | (data.col >= num_dims)).sum())   # int32 bit patterns of uint32 ids
        if bad and rank == 0:
            print(f"[node {engine.node.id}] {bad} feature ids of {ctx.get_string('input')} are outside "
                  f"[0, {num_dims}) (num_dims) and are ignored, like the reference's Slice drops them")
    else:
        data = make_synthetic(ctx.get_int32("num_samples"), num_dims, ctx.get_int32("nnz_per_row"), dev,
                              seed=ctx.get_int32("seed") + rank)

    if not ctx.get_bool("keep_empty_nodes"):
        # lr_example.cpp:145-152: a node that read no data leaves the cluster before the tables exist; the
        # others go on without it (its GPU holds no shard)
        if not engine.force_quit(data.n_rows == 0):
            print(f"[node {engine.node.id}] ForceQuit the engine as no data allocated", flush=True)
            engine.stop_everything()
            return
        rank, world = engine.rank_index, len(engine.nodes)
    model = ModelType[ctx.get_string("kModelType")]
    storage = StorageType[ctx.get_string("kStorageType")]
    dtype = torch.float64 if ctx.get_string("dtype") == "fp64" else torch.float32
    tid = engine.create_table(num_dims=num_dims, model_type=model, storage_type=storage,
                              staleness=ctx.get_int32("kStaleness"), dtype=dtype)
    if resume:
        engine.context["checkpoint_file_prefix"] = ctx.get_string("checkpoint_file_prefix")
        its = engine.restore_table(tid)
        print(f"[node {engine.node.id}] resumed from checkpoint at iteration {min(its.values()) if its else 0}",
              flush=True)
    engine.barrier()

    B = ctx.get_int32("batch_size")
    iters = ctx.get_int32("num_iters")
    alpha = ctx.get_double("alpha")
    ck_int = ctx.get_int32("checkpoint_interval")
    ck_on = ctx.get_bool("checkpoint_toggle")
    rep_int = ctx.get_int32("report_interval")
    rep_prefix = ctx.get_string("report_prefix")
    p_straggle = ctx.get_double("with_injected_straggler")
    pause_at, pause_s = ctx.get_int32("pause_at_iter"), ctx.get_double("pause_s")
    workers_per_node = max(1, ctx.get_int32("num_workers_per_node"))
    # --keep_empty_nodes: a node without local samples stays as a server-only rank (its HBM keeps its share of
    # the table); it takes part in every collective of the run but runs no worker
    has_data = [bool(x) for x in engine.control.all_gather(data.n_rows > 0)]
    if not has_data[rank]:
        print(f"[node {engine.node.id}] no local data: serving its shards only", flush=True)
    result = {}

    def worker(info):
        client = info.create_kv_client_table(tid)
        tr = SparseLRTrainer(client, data, alpha)
        g = torch.Generator().manual_seed(ctx.get_int32("seed") * 1000 + info.worker_id)
        starts = torch.randint(0, max(1, data.n_rows - B + 1), (iters,), generator=g)  # precomputed like :231-235
        dense = tr.dense_regime(B) and workers_per_node == 1
        first = engine.get_iteration(info.worker_id) if resume else 0
        t0 = time.time()
        i = first
        while i < iters:
            if job is not None:
                job.check()
            # run up to the next checkpoint / report / pause boundary inside one persistent kernel
            stop = iters
            for step in (ck_int if ck_on else 0, rep_int):
                if step and step > 0:
                    stop = min(stop, (i // step + 1) * step)
            if pause_at > i:
                stop = min(stop, pause_at)
            if dense:
                tr.train_dense(starts[i:stop].to(dev), B)
            else:
                for j in range(i, stop):
                    tr.step_sparse(int(starts[j]), B)
            i = stop
            engine.set_iteration(info.worker_id, i)
            if ck_on and ck_int > 0 and i % ck_int == 0:
                c0 = time.time()
                client.checkpoint()  # collective on this backend
                if info.worker_id == 0:
                    print(f"[worker 0] checkpoint at iteration {i} took {(time.time() - c0) * 1e3:.1f} ms", flush=True)
            if rep_int > 0 and i % rep_int == 0 and rep_prefix and info.worker_id == 0:
                torch.cuda.synchronize()
                with open(f"{rep_prefix}_{engine.node.id}", "a") as f:
                    f.write(f"{i}\t{(time.time() - t0) * 1e3:.3f}\n")
            if p_straggle > 0 and random.random() < p_straggle:
                # lr_example.cpp:347-353 sleeps the worker thread; here the worker's STREAM is held, so its
                # next clock is late on the device timeline and peers meet the consistency gate for real
                client.straggle(random.random() * 0.1)
            if pause_at > 0 and i == pause_at and not resume:
                # fault-injection aid: hold the run at a known iteration (the kill tests aim here)
                torch.cuda.synchronize()
                if rep_prefix:
                    open(f"{rep_prefix}_paused_{engine.node.id}", "w").write(str(i))
                if job is not None:
                    job.sleep(pause_s)
                else:
                    time.sleep(pause_s)
        torch.cuda.synchronize()
        elapsed = (time.time() - t0) * 1e3
        loss, acc = tr.evaluate()
        result[info.worker_id] = (elapsed, loss, acc)
        print(f"[node {engine.node.id} worker {info.worker_id}] Total training time: {elapsed:.1f} ms, "
              f"logloss {loss:.4f}, accuracy {acc:.4f} on {data.n_rows} local samples", flush=True)

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(n.id, workers_per_node) for n, ok in zip(engine.nodes, has_data) if ok])
    task.set_tables([tid])
    task.set_lambda(worker)
    engine.run(task)
    wout = ctx.get_string("weights_out")
    if wout:
        torch.save(engine.tables[tid].gather_full().cpu(), f"{wout}_{engine.node.id}.pt")
    engine.stop_everything()


def main(argv=None):
    from ..parallel.engine import Engine
    from ..parallel.types import Node
    from ..utils.flags import Context, define_lr_flags

    ctx = Context.get_instance()
    define_lr_flags(ctx)
    ctx.define("dtype", "fp32", "fp32 | fp64: value type of the table (the reference runs double)")
    ctx.define("store_port", 29700, "rendezvous store of the fault-tolerant mode (hosted by the master)")
    ctx.define("relaunched", False, "set by the master's relaunch hook: wait for its rollback order before joining")
    ctx.define("weights_out", "", "write the final table to <prefix>_<node id>.pt")
    ctx.define("keep_empty_nodes", False, "a node without data serves its shards instead of leaving (ForceQuit)")
    ctx.define("pause_at_iter", 0, "fault-injection aid: hold the first generation at this iteration")
    ctx.define("pause_s", 0.0, "... for this many seconds")
    ctx.parse(argv)
    hb = ctx.get_int32("heartbeat_interval")
    if hb > 0 and ctx.get_string("config_file"):
        # fault-tolerant mode: master = node id 1 (python -m minips_b200.apps.master), this process is one
        # of the other nodes of the hostfile; one generation per rollback (parallel/fault.py)
        from .. import core
        from ..parallel.fault import FaultTolerantJob
        os.environ.setdefault("MINIPS_SPIN_TIMEOUT_S", "5")
        core.context_set("checkpoint_file_prefix", ctx.get_string("checkpoint_file_prefix"))
        scale_node = None
        if ctx.get_bool("scale"):
            # elastic scale-out (lr_example.cpp:396-409): this process is a NEW node joining a running job
            scale_node = core.Node(ctx.get_int32("my_id"), ctx.get_string("scale_name") or "localhost",
                                   ctx.get_int32("scale_port"))
            print(f"[node {scale_node.id}] scaling out: joining as {scale_node.hostname}:{scale_node.port}", flush=True)
        job = FaultTolerantJob(core.parse_file(ctx.get_string("config_file")), ctx.get_int32("my_id"), hb,
                               ctx.get_int32("store_port"), relaunched=ctx.get_bool("relaunched"),
                               scale_node=scale_node, scale_file=ctx.get_string("scale_file"))
        job.start()
        gpu = {}

        def body(job):
            if "device" not in gpu:       # a process keeps its GPU across generations
                gpu["device"] = int(os.environ.get("LOCAL_RANK", job.rank))
                torch.cuda.set_device(gpu["device"])
            device = gpu["device"]
            nodes = [Node(n.id, n.hostname, n.port) for n in job.rank_nodes]
            engine = Engine(nodes[job.rank], nodes, device=device)
            try:
                _train(engine, ctx, job)
            finally:
                try:
                    engine.stop_everything()
                except Exception:  # noqa: BLE001 - a generation that failed leaves half-open peers behind
                    pass

        job.run(body)
        job.stop()
        return 0
    engine = Engine.from_env()
    ctx.set("my_id", engine.node.id)
    _train(engine, ctx)
    return 0


if __name__ == "__main__":
    sys.exit(main())
