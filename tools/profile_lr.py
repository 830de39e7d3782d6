"""Per-iteration time of the dense LR kernel vs batch size / consistency model (1 GPU),
its in-kernel phase timeline, or a short fixed run for ncu (--ncu)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minips_b200.models.lr import SparseLRTrainer, make_synthetic  # noqa: E402
from minips_b200.parallel.engine import Engine  # noqa: E402
from minips_b200.parallel.types import MLTask, ModelType, Node, WorkerAlloc  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--ncu", action="store_true")
ap.add_argument("--rows", type=int, default=2_000_000)
ap.add_argument("--features", type=int, default=1_000_000)
ap.add_argument("--dtype", default="fp64", choices=["fp64", "fp32"])
a = ap.parse_args()
DTYPE = torch.float64 if a.dtype == "fp64" else torch.float32

node = Node(0, "localhost", 0)
engine = Engine(node, [node], device=0)
engine.start_everything(num_server_threads_per_node=8)
data = make_synthetic(a.rows, a.features, 40, seed=1)
g = torch.Generator().manual_seed(3)
PHASES = ["gate", "pull", "sync1", "rows+sync2", "pulled-wait", "push+signal"]


def run(model, batch, iters, timeline=False):
    tid = engine.create_table(num_dims=a.features, model_type=model, staleness=2, dtype=DTYPE)
    res = {}

    def fn(info):
        tr = SparseLRTrainer(info.create_kv_client_table(tid), data, 0.01)
        st = torch.randint(0, a.rows - batch, (iters,), generator=g).to("cuda")
        tr.train_dense(st[:3].contiguous(), batch)
        torch.cuda.synchronize()
        tl = torch.zeros(iters * 8, dtype=torch.int64, device="cuda") if timeline else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tr.train_dense(st, batch, timeline=tl)
        e1.record()
        torch.cuda.synchronize()
        res["us"] = e0.elapsed_time(e1) * 1e3 / iters
        if timeline:
            t = tl.view(iters, 8).cpu().double()
            d = (t[:, 1:6] - t[:, 0:5]) / 1e3
            nxt = (t[1:, 0] - t[:-1, 5]) / 1e3
            res["phases"] = d[2:].mean(0).tolist() + [nxt[2:].mean().item(),
                                                      ((t[:, 6] - t[:, 4]) / 1e3)[2:].mean().item()]

    t = MLTask()
    t.set_worker_alloc([WorkerAlloc(0, 1)])
    t.set_tables([tid])
    t.set_lambda(fn)
    engine.run(t)
    engine.tables[tid].close()
    return res


if a.ncu:
    print("ncu run us/iter", run(ModelType.BSP, 65536, 2)["us"])
else:
    for model in (ModelType.BSP, ModelType.ASP):
        for batch in (1, 4096, 16384, 65536, 262144):
            us = run(model, batch, 50)["us"]
            print(f"{model.name} batch={batch:7d}  {us:9.2f} us/iter  {batch / us:9.1f} Msamples/s", flush=True)
    for batch in (1, 65536):
        r = run(ModelType.BSP, batch, 30, timeline=True)
        names = ["gate", "pull+sync1", "rows+sync2", "pulled-wait", "push+signal", "gap-to-next-iter", "push-loop-only"]
        print(f"BSP batch={batch} phases us (CTA0 thread0):",
              ", ".join(f"{n}={v:.2f}" for n, v in zip(names, r["phases"])), f"total={r['us']:.2f}")
engine.stop_everything()
