"""BasicExample — the smoke application of the reference (apps/basic/basic_example.cpp:25-76):
a 1000-key table under SSP(staleness 1) with Map storage, 10 workers per node, each doing
100 x (Get all keys, Add 0.5 to every key, Clock).  Afterwards every key holds
0.5 * 100 * #workers.

    python -m minips_b200.apps.basic_example              # one GPU
    torchrun --nproc-per-node 2 -m minips_b200.apps.basic_example
"""
from __future__ import annotations

import sys

import torch


def main(argv=None):
    from ..parallel.engine import Engine
    from ..parallel.types import MLTask, ModelType, StorageType, WorkerAlloc
    from ..utils.flags import Context, define_common_flags

    ctx = Context.get_instance()
    define_common_flags(ctx)
    ctx.define("num_keys", 1000)
    ctx.define("workers", 10)
    ctx.define("iters", 100)
    ctx.parse(argv)
    engine = Engine.from_env()
    world = len(engine.nodes)
    engine.start_everything()
    n_keys, W, iters = ctx.get_int32("num_keys"), ctx.get_int32("workers"), ctx.get_int32("iters")
    per = n_keys // world
    ranges = [(i * per, (i + 1) * per if i < world - 1 else n_keys) for i in range(world)]
    tid = engine.create_table(ranges=ranges, model_type=ModelType.SSP, storage_type=StorageType.Map, staleness=1)
    engine.barrier()
    keys = torch.arange(n_keys)

    def worker(info):
        table = info.create_kv_client_table(tid)
        half = torch.full((n_keys,), 0.5)
        for _ in range(iters):
            table.get(keys)
            table.add(keys, half)
            table.clock()

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(n.id, W) for n in engine.nodes])
    task.set_tables([tid])
    task.set_lambda(worker)
    engine.run(task)
    full = engine.tables[tid].gather_full().cpu()
    expect = 0.5 * iters * W * world
    ok = bool(torch.all(full == expect))
    if engine.node.id == 0:
        print(f"BasicExample: every key == {expect}: {ok}")
    engine.stop_everything()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
