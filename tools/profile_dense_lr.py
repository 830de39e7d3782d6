"""Throughput of the dense-LR path: tcgen05 forward with the fused weight pull and the
sigmoid/loss epilogue, the gradient kernel, and the whole Get -> forward -> grad -> Add -> Clock
iteration, against plain PyTorch (bf16 matmul + sigmoid + matmul) on the same inputs.

The op is a matrix-vector product (one weight vector), so the roofline is HBM bandwidth: a forward reads
the batch once (n x D bf16), an iteration reads it twice (forward + gradient).
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minips_b200.models.dense_lr import DenseLRTrainer  # noqa: E402
from minips_b200.parallel.engine import Engine  # noqa: E402
from minips_b200.parallel.types import MLTask, ModelType, Node, WorkerAlloc  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=2_000_000)
ap.add_argument("--D", type=int, default=1024)
ap.add_argument("--batch", type=int, default=1 << 20)
ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()

HBM_GBS = 6448.7
try:
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
        HBM_GBS = float(json.load(f).get("hbm_gbs", HBM_GBS))
except (OSError, ValueError):
    pass

node = Node(0, "localhost", 0)
engine = Engine(node, [node], device=0)
engine.start_everything(num_server_threads_per_node=8)
g = torch.Generator(device="cuda").manual_seed(0)
x = (torch.randn(a.n, a.D, device="cuda", generator=g) * 0.1).to(torch.bfloat16)
w_true = torch.randn(a.D, device="cuda", generator=g)
y = ((x.float() @ w_true) > 0).float()
tid = engine.create_table(num_dims=a.D, model_type=ModelType.BSP)
B = min(a.batch, a.n)
starts = [int(s) for s in torch.randint(0, max(1, a.n - B), (a.iters + 3,))]
out = {}


def timed(fn, n):
    for s in starts[:3]:
        fn(s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in starts[3:3 + n]:
        fn(s)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def worker(info):
    tr = DenseLRTrainer(info.create_kv_client_table(tid), x, y, alpha=1e-3 / B)
    out["mode"] = tr.mode
    out["forward_ms"] = timed(lambda s: tr.forward(s, B), a.iters)
    out["step_ms"] = timed(lambda s: tr.step(s, B), a.iters)
    w = torch.zeros(a.D, device="cuda")

    def torch_step(s):
        xb = x[s:s + B]
        p = torch.sigmoid(xb @ w.to(torch.bfloat16)).float()
        coef = (1e-3 / B) * (y[s:s + B] - p)
        w.add_((coef.to(torch.bfloat16) @ xb).float())

    out["torch_step_ms"] = timed(torch_step, a.iters)
    out["torch_forward_ms"] = timed(lambda s: torch.sigmoid(x[s:s + B] @ w.to(torch.bfloat16)), a.iters)


task = MLTask()
task.set_worker_alloc([WorkerAlloc(0, 1)])
task.set_tables([tid])
task.set_lambda(worker)
engine.run(task)
engine.stop_everything()
gb = B * a.D * 2 / 1e9
print(f"dense LR, batch {B} x D {a.D} bf16 ({gb:.2f} GB per pass), HBM copy peak {HBM_GBS:.0f} GB/s")
print(f"  forward (tcgen05, pull + sigmoid/loss fused): {out['forward_ms']:.3f} ms  "
      f"{gb / out['forward_ms'] * 1e3:.0f} GB/s"
      f"  = {gb / out['forward_ms'] * 1e3 / HBM_GBS:.2f} of HBM copy peak   "
      f"[torch matmul+sigmoid: {out['torch_forward_ms']:.3f} ms]")
passes = 1 if out["mode"] in ("rows", "stream", "mma") else 2   # how often the iteration reads X from HBM
print(f"  iteration [{out['mode']}: X read {passes}x] (Get, forward, gradient, Add, Clock): {out['step_ms']:.3f} ms  "
      f"{passes * gb / out['step_ms'] * 1e3:.0f} GB/s of X traffic"
      f"  = {passes * gb / out['step_ms'] * 1e3 / HBM_GBS:.2f} of HBM copy peak   "
      f"{B / out['step_ms'] / 1e6:.2f} G samples/s"
      f"   [torch: {out['torch_step_ms']:.3f} ms, {B / out['torch_step_ms'] / 1e6:.2f} G samples/s]")
