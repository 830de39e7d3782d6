"""Push / pull micro-benchmark (BASELINE.json config 5): keys/s and GB/s vs batch size
(1K - 10M keys) for scattered keys (mps_pull / mps_push) and for contiguous whole-table slabs
(mps_pull_dense / mps_push_dense), device-timed, max over ranks.

    python tools/bench_push_pull.py                       # 1 GPU, 8 local shards
    torchrun --nproc-per-node 8 tools/bench_push_pull.py  # 8 GPUs, 1 shard each

With N ranks a fraction (N-1)/N of the keys lives on peers; the NVLink roofline for that
traffic is (remote payload bytes) / 770 GB/s per direction per GPU (measured peer-copy
bandwidth, B200_PROFILING.md); the achieved fraction is printed per row.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minips_b200.parallel.engine import Engine  # noqa: E402
from minips_b200.parallel.types import MLTask, ModelType, WorkerAlloc  # noqa: E402
from minips_b200.utils.benchkit import max_over_ranks, sync_barrier  # noqa: E402

NVLINK_GBS = 770.0


def main():
    engine = Engine.from_env()
    world, rank = len(engine.nodes), engine.node.id
    dev = f"cuda:{engine.device}"
    engine.start_everything(num_server_threads_per_node=max(1, 8 // world))
    num_dims = int(os.environ.get("MPS_MICRO_DIMS", 100_000_000))
    tid = engine.create_table(num_dims=num_dims, model_type=ModelType.ASP)
    rows = []

    def timed(fn, iters=10, warm=3):
        for _ in range(warm):
            fn()
        sync_barrier(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        sync_barrier(dev)
        return max_over_ranks(e0.elapsed_time(e1) / iters, dev)

    def worker(info):
        t = info.create_kv_client_table(tid)
        g = torch.Generator(device=dev).manual_seed(5 + rank)
        for n in (1_000, 10_000, 100_000, 1_000_000, 10_000_000):
            keys = torch.sort(torch.randint(0, num_dims, (n,), device=dev, generator=g))[0].to(torch.int32)
            vals = torch.ones(n, device=dev)
            out = torch.empty(n, device=dev)
            ms_pull = timed(lambda: t.get(keys, out=out))
            ms_push = timed(lambda: t.add(keys, vals))
            remote = (world - 1) / world
            for op, ms in (("pull", ms_pull), ("push", ms_push)):
                payload = n * 4
                roof_ms = payload * remote / (NVLINK_GBS * 1e6) if world > 1 else None
                rows.append({"op": op, "pattern": "scattered", "keys": n, "ms": ms, "Mkeys_per_s": n / ms / 1e3,
                             "payload_GBps": payload / ms / 1e6,
                             "nvlink_roofline_frac": (roof_ms / ms) if roof_ms else None})
        # contiguous: the whole table as slabs
        full = torch.empty(num_dims, device=dev)
        delta = torch.ones(num_dims, device=dev)
        ms_pull = timed(lambda: t.get_all(out=full), iters=5)
        ms_push = timed(lambda: t.add_all(delta), iters=5)
        remote = (world - 1) / world
        for op, ms in (("pull", ms_pull), ("push", ms_push)):
            payload = num_dims * 4
            roof_ms = payload * remote / (NVLINK_GBS * 1e6) if world > 1 else None
            rows.append({"op": op, "pattern": "contiguous", "keys": num_dims, "ms": ms,
                "Mkeys_per_s": num_dims / ms / 1e3,
                         "payload_GBps": payload / ms / 1e6,
                         "nvlink_roofline_frac": (roof_ms / ms) if roof_ms else None})

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(n.id, 1) for n in engine.nodes])
    task.set_tables([tid])
    task.set_lambda(worker)
    engine.run(task)
    engine.stop_everything()
    if rank == 0:
        print(json.dumps({"bench": "push_pull_micro", "n_gpus": world, "table_keys": num_dims, "rows": rows}))
        for r in rows:
            frac = f"{r['nvlink_roofline_frac']:.2f}" if r["nvlink_roofline_frac"] is not None else "  - "
            print(f"{r['op']:5s} {r['pattern']:10s} keys={r['keys']:>10d}  {r['ms']:9.4f} ms  "
                  f"{r['Mkeys_per_s']:10.1f} Mkeys/s"
                  f"  {r['payload_GBps']:8.1f} GB/s  roofline {frac}", file=sys.stderr)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
