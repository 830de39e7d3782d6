"""Throughput of the tcgen05 k-means assignment kernel vs the CUDA-core kernel and
torch (cuBLAS bf16 matmul + argmin, which materialises the n x K distance matrix)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minips_b200.models.kmeans import KMeansAssigner  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=4_000_000)
ap.add_argument("--K", type=int, default=1024)
ap.add_argument("--D", type=int, default=128)
ap.add_argument("--ncu", action="store_true")
a = ap.parse_args()

g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn(a.n, a.D, device="cuda", generator=g).to(torch.bfloat16)
centers = torch.randn(a.K, a.D, device="cuda", generator=g)
km = KMeansAssigner(a.K, a.D, "cuda")
out = (torch.empty(a.n, dtype=torch.int32, device="cuda"), torch.empty(a.n, dtype=torch.float32, device="cuda"))
flops = 2.0 * a.n * a.K * a.D


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


if a.ncu:
    km.assign(x, centers=centers, out=out)
    km.assign(x, centers=centers, out=out)
    torch.cuda.synchronize()
    print("ncu run done")
    sys.exit(0)

ms = timed(lambda: km.assign(x, centers=centers, out=out))
print(f"tcgen05 assign   n={a.n} K={a.K} D={a.D}: {ms:8.3f} ms  {flops / ms / 1e9:8.1f} TFLOP/s  "
      f"{a.n / ms / 1e3:8.1f} Mpoints/s")

nr = min(a.n, 200_000)
ms_r = timed(lambda: km.assign_ref(x[:nr]), iters=2, warm=1)
print(f"cuda-core kernel n={nr}: {ms_r:8.3f} ms  {2.0 * nr * a.K * a.D / ms_r / 1e9:8.1f} TFLOP/s")

cb = centers.to(torch.bfloat16)
cn = (cb.float() ** 2).sum(1)
chunk = 1 << 20


def torch_assign():
    for s in range(0, a.n, chunk):
        d = torch.addmm(cn.to(torch.bfloat16)[None, :], x[s:s + chunk], cb.T, alpha=-2.0)
        out[0][s:s + chunk] = d.argmin(1).int()


ms_t = timed(torch_assign, iters=3, warm=1)
print(f"torch addmm+argmin (bf16, 1M-row chunks): {ms_t:8.3f} ms  {flops / ms_t / 1e9:8.1f} TFLOP/s")
