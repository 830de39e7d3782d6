#!/usr/bin/env python
"""NVLink store / fence micro-measurements (csrc/kernels/probe.cu) on N >= 2 GPUs:
    torchrun --nproc-per-node 2 tools/nvlink_probe.py
Rank 0 stores `bytes` into rank 1's buffer (16-byte st, 16-byte red, copy-engine bulk store), into its
own buffer, or into the multicast mapping, and reports when the stores were issued and when a
fence.acq_rel.sys behind them returned (= when a flag could be published)."""
import ctypes as C
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minips_b200.ops import _lib  # noqa: E402
from minips_b200.parallel.engine import Engine  # noqa: E402
from minips_b200.parallel.heap import SymmetricBuffer  # noqa: E402


def main():
    engine = Engine.from_env()
    engine.start_everything()
    rank, world = engine.node.id, len(engine.nodes)
    dev = f"cuda:{engine.device}"
    lib = _lib.load()
    lib.mps_probe_store.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    lib.mps_probe_store.restype = C.c_int
    cap = 16 << 20
    buf = SymmetricBuffer(cap, engine.control, engine.device, multicast=True)
    src = torch.randn(cap // 4, device=dev)
    out = torch.zeros(3, dtype=torch.int64, device=dev)
    engine.barrier()
    if rank == 0:
        peer = buf.ptrs[1 % world]
        dsts = [("peer", peer), ("local", buf.ptrs[0])] + ([("multicast", buf.mc_ptr)] if buf.mc_ptr else [])
        modes = [("st.v4", 0), ("multimem.st.v4", 1), ("bulk 16K", 2), ("red.v4", 3)]
        print(f"world={world} multicast={'yes' if buf.mc_ptr else 'no'}; ns = max over CTAs; GB/s = bytes / fence-done")
        print(f"{'dst':10s} {'op':15s} {'bytes':>9s} {'issued':>8s} {'fence-done':>10s} {'fence':>7s} {'GB/s':>7s}")
        for dname, dptr in dsts:
            for mname, m in modes:
                if (m == 1) != (dname == "multicast"):
                    continue
                for nbytes in (0, 1 << 19, 1 << 20, 2 << 20, 4 << 20, 8 << 20):
                    rows = []
                    for i in range(25):
                        _lib.check(lib.mps_probe_store(m, C.c_void_p(dptr), C.c_void_p(src.data_ptr()), nbytes,
                                                       C.c_void_p(out.data_ptr()), _lib.stream_ptr()), "probe")
                        torch.cuda.synchronize()
                        if i >= 5:
                            rows.append(out.tolist())
                    med = [statistics.median(r[k] for r in rows) for k in range(3)]
                    gbs = nbytes / med[1] if med[1] else 0.0
                    print(f"{dname:10s} {mname:15s} {nbytes:9d} {med[0]:8.0f} {med[1]:10.0f} {med[2]:7.0f} {gbs:7.1f}")
    engine.barrier()
    buf.close()
    engine.stop_everything()
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
