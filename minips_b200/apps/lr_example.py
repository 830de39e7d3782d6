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

import random
import sys
import time

import torch


def main(argv=None):
    from ..models.lr import SparseLRTrainer, make_synthetic
    from ..ops.sparse import load_libsvm_file
    from ..parallel.engine import Engine
    from ..parallel.types import MLTask, ModelType, StorageType, WorkerAlloc
    from ..utils.flags import Context, define_lr_flags

    ctx = Context.get_instance()
    define_lr_flags(ctx)
    ctx.parse(argv)
    engine = Engine.from_env()
    rank, world = engine.node.id, len(engine.nodes)
    ctx.set("my_id", rank)
    engine.context.update(ctx.values)
    dev = f"cuda:{engine.device}"
    servers = max(1, ctx.get_int32("num_servers_per_node"))
    engine.start_everything(num_server_threads_per_node=servers)

    num_dims = ctx.get_int32("num_dims")
    if ctx.get_string("input"):
        data = load_libsvm_file(ctx.get_string("input"), dev)
        # every node parses the file and keeps its contiguous share of the rows
        per = data.n_rows // world
        data = data.rows(rank * per, per if rank < world - 1 else data.n_rows - rank * per)
        data.col = data.col.contiguous()
        data.val = data.val.contiguous()
    else:
        data = make_synthetic(ctx.get_int32("num_samples"), num_dims, ctx.get_int32("nnz_per_row"), dev,
                              seed=ctx.get_int32("seed") + rank)
    if data.n_rows == 0:
        print(f"[node {rank}] no local data: leaving the cluster (ForceQuit)")  # lr_example.cpp:145-152
        engine.stop_everything()
        return 0

    model = ModelType[ctx.get_string("kModelType")]
    storage = StorageType[ctx.get_string("kStorageType")]
    tid = engine.create_table(num_dims=num_dims, model_type=model, storage_type=storage,
                              staleness=ctx.get_int32("kStaleness"))
    if ctx.get_bool("use_weight_file"):
        engine.context["checkpoint_file_prefix"] = ctx.get_string("checkpoint_file_prefix")
        its = engine.restore_table(tid)
        print(f"[node {rank}] resumed from checkpoint at iteration {min(its.values()) if its else 0}")
    engine.barrier()

    B = ctx.get_int32("batch_size")
    iters = ctx.get_int32("num_iters")
    alpha = ctx.get_double("alpha")
    ck_int = ctx.get_int32("checkpoint_interval")
    ck_on = ctx.get_bool("checkpoint_toggle")
    rep_int = ctx.get_int32("report_interval")
    rep_prefix = ctx.get_string("report_prefix")
    p_straggle = ctx.get_double("with_injected_straggler")
    workers_per_node = max(1, ctx.get_int32("num_workers_per_node"))
    result = {}

    def worker(info):
        client = info.create_kv_client_table(tid)
        tr = SparseLRTrainer(client, data, alpha)
        g = torch.Generator().manual_seed(ctx.get_int32("seed") * 1000 + info.worker_id)
        starts = torch.randint(0, max(1, data.n_rows - B + 1), (iters,), generator=g)  # precomputed like :231-235
        dense = tr.dense_regime(B) and workers_per_node == 1
        first = engine.get_iteration(info.worker_id) if ctx.get_bool("use_weight_file") else 0
        t0 = time.time()
        i = first
        while i < iters:
            # run up to the next checkpoint / report boundary inside one persistent kernel
            stop = iters
            for step in (ck_int if ck_on else 0, rep_int):
                if step and step > 0:
                    stop = min(stop, (i // step + 1) * step)
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
                    print(f"[worker 0] checkpoint at iteration {i} took {(time.time() - c0) * 1e3:.1f} ms")
            if rep_int > 0 and i % rep_int == 0 and rep_prefix and info.worker_id == 0:
                torch.cuda.synchronize()
                with open(f"{rep_prefix}_{rank}", "a") as f:
                    f.write(f"{i}\t{(time.time() - t0) * 1e3:.3f}\n")
            if p_straggle > 0 and random.random() < p_straggle:
                # lr_example.cpp:347-353 sleeps the worker thread; here the worker's STREAM is held, so its
                # next clock is late on the device timeline and peers meet the consistency gate for real
                client.straggle(random.random() * 0.1)
        torch.cuda.synchronize()
        elapsed = (time.time() - t0) * 1e3
        loss, acc = tr.evaluate()
        result[info.worker_id] = (elapsed, loss, acc)
        print(f"[node {rank} worker {info.worker_id}] Total training time: {elapsed:.1f} ms, "
              f"logloss {loss:.4f}, accuracy {acc:.4f} on {data.n_rows} local samples")

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(n.id, workers_per_node) for n in engine.nodes])
    task.set_tables([tid])
    task.set_lambda(worker)
    engine.run(task)
    engine.stop_everything()
    return 0


if __name__ == "__main__":
    sys.exit(main())
