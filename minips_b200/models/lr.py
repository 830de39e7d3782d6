"""Sparse logistic regression on the parameter server (reference apps/logistic_regression/
lr_example.cpp:97-385): mini-batch SGD where every iteration is Get(batch keys) -> forward ->
gradient -> Add -> Clock.  The device implementation runs those as phases of one fused
kernel (csrc/kernels/lr.cu); this module holds the data container, the synthetic libsvm-shaped
generator that replaces the HDFS loader, the trainer and a plain PyTorch fp32 oracle.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from ..ops import _lib
from ..parallel.device_table import DeviceTable, KVClientTable
from ..parallel.types import ModelType
from ..utils.tracing import traced


@dataclass
class CsrBatchSource:
    """Local samples of one node in CSR form (labels in {0,1}; reference maps y<0 to 0,
    lr_example.cpp:296-297; feature ids 0-based, lib/parser.hpp:33)."""
    row_ptr: torch.Tensor  # int64 [n+1]
    col: torch.Tensor      # int32 [nnz] (uint32 bit patterns)
    val: torch.Tensor      # float32 [nnz]
    label: torch.Tensor    # float32 [n]

    @property
    def n_rows(self) -> int:
        return self.label.numel()

    @property
    def nnz(self) -> int:
        return self.col.numel()

    def to(self, device) -> "CsrBatchSource":
        return CsrBatchSource(self.row_ptr.to(device), self.col.to(device), self.val.to(device),
                              self.label.to(device))

    def pin(self) -> "CsrBatchSource":
        return CsrBatchSource(self.row_ptr.pin_memory(), self.col.pin_memory(), self.val.pin_memory(),
                              self.label.pin_memory())

    def c_struct(self) -> _lib.Csr:
        m = _lib.Csr()
        m.row_ptr = self.row_ptr.data_ptr()
        m.col = self.col.data_ptr()
        m.val = self.val.data_ptr()
        m.label = self.label.data_ptr()
        m.n_rows = self.n_rows
        return m

    def rows(self, start: int, count: int) -> "CsrBatchSource":
        """Contiguous row slice [start, start+count) (no wrap) re-based to its own row_ptr."""
        rp = self.row_ptr[start:start + count + 1]
        b, e = int(rp[0]), int(rp[-1])
        return CsrBatchSource(rp - rp[0], self.col[b:e], self.val[b:e], self.label[start:start + count])


def make_synthetic(n_rows: int, n_features: int, nnz_per_row: int = 40, device="cuda", seed: int = 0,
                   binary_values: bool = False, chunk_rows: int = 1 << 20) -> CsrBatchSource:
    """libsvm-shaped synthetic data generated directly in device memory: uniform feature ids,
    N(0,1) values, labels drawn from a hidden linear model."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    w_true = torch.randn(n_features, device=device, generator=torch.Generator(device=device).manual_seed(12345))
    col = torch.empty(n_rows * nnz_per_row, dtype=torch.int32, device=device)
    val = torch.empty(n_rows * nnz_per_row, dtype=torch.float32, device=device)
    label = torch.empty(n_rows, dtype=torch.float32, device=device)
    for r0 in range(0, n_rows, chunk_rows):
        n = min(chunk_rows, n_rows - r0)
        c = torch.randint(0, n_features, (n, nnz_per_row), device=device, generator=g, dtype=torch.int64)
        c, _ = torch.sort(c, dim=1)
        v = torch.ones(n, nnz_per_row, device=device) if binary_values else \
            torch.randn(n, nnz_per_row, device=device, generator=g)
        z = (w_true[c] * v).sum(1) * (4.0 / nnz_per_row ** 0.5)
        y = (torch.rand(n, device=device, generator=g) < torch.sigmoid(z)).float()
        col[r0 * nnz_per_row:(r0 + n) * nnz_per_row] = c.reshape(-1).to(torch.int32)
        val[r0 * nnz_per_row:(r0 + n) * nnz_per_row] = v.reshape(-1)
        label[r0:r0 + n] = y
    row_ptr = torch.arange(0, (n_rows + 1) * nnz_per_row, nnz_per_row, dtype=torch.int64, device=device)
    return CsrBatchSource(row_ptr, col, val, label)


class XchgScratch:
    """Symmetric scratch of the reduce-at-owner / owner-broadcasts kernel (csrc/kernels/lr_xchg.cu):
    per rank one weight replica and two gradient accumulators, mapped by every peer and — when the
    fabric has NVLS — by the switch as one multicast address.  Creating it is collective."""

    def __init__(self, table: DeviceTable):
        import os
        from ..parallel.heap import SymmetricBuffer
        n = table.num_keys
        self.stride_bytes = (n * table.itemsize + 255) // 256 * 256
        self.buf = SymmetricBuffer(3 * self.stride_bytes, table.control, table.device, multicast=True)
        x = _lib.Xchg()
        for r in range(table.world):
            x.replica[r] = self.buf.ptrs[r]
            x.gacc[r] = self.buf.ptrs[r] + self.stride_bytes
        x.replica_mc = self.buf.mc_ptr or None
        x.gacc_mc = (self.buf.mc_ptr + self.stride_bytes) if self.buf.mc_ptr else None
        x.gacc_stride = self.stride_bytes // table.itemsize
        # Measured on B200 / NVSwitch (tools/xchg_sweep.sh, profiles/round2/): fp32 is fastest through NVLS
        # (multimem.ld_reduce.v4 / multimem.st.v4) from 4 GPUs up; fp64 has only a scalar
        # multimem.ld_reduce and its 8-byte multicast stores drain slowly, so 16-byte peer loads / stores
        # win at every size; with two GPUs a plain peer store beats the multicast store for fp32 too.
        how_r = os.environ.get("MINIPS_XCHG_REDUCE", "auto")
        how_b = os.environ.get("MINIPS_XCHG_BCAST", "auto")
        f64 = table.dtype == torch.float64
        p2p_r = how_r == "p2p" or (how_r == "auto" and f64)
        p2p_b = how_b == "p2p" or (how_b == "auto" and (f64 or table.world <= 2))
        x.flags = (_lib.XCHG_REDUCE_P2P if p2p_r else 0) | (_lib.XCHG_BCAST_P2P if p2p_b else 0)
        self.c = x
        self.multicast = bool(self.buf.mc_ptr)
        self.reduce = "p2p" if (p2p_r or not self.multicast) else "multimem"
        self.bcast = "p2p" if (p2p_b or not self.multicast) else "multimem"
        self.launch_seq = 0         # exchange launches so far (picks the counter set; same on every rank)
        self.valid_clock = -1       # clock whose weights the local replica holds
        self.valid_epoch = -1       # table.write_epoch at that time
        self.replica = self.buf.local_tensor(0, n, table.dtype)

    def close(self):
        self.buf.close()


class BucketScratch:
    """Scratch of the bucketed (message-flow) sparse step, csrc/kernels/lr_bucket.cu: per rank an inbox for
    the keys and deltas of every requester, a reply buffer for its own requests (all peer-mapped) and local
    staging.  cap = the largest number of non-zeros a batch can have.  Creating it is collective."""

    def __init__(self, table: DeviceTable, cap: int):
        import os
        from ..parallel.heap import SymmetricBuffer
        R, it = table.world, table.itemsize
        self.cap = cap = (int(cap) + 63) // 64 * 64
        a256 = lambda n: (n + 255) // 256 * 256  # noqa: E731
        off_keys = 0
        off_delta = off_keys + a256(R * cap * 4)
        off_meta = off_delta + a256(R * cap * it)
        off_reply = off_meta + a256(R * 2 * 8)
        total = off_reply + a256(cap * it)
        self.buf = SymmetricBuffer(total, table.control, table.device)
        dev = f"cuda:{table.device}"
        self.send_keys = torch.empty(cap, dtype=torch.int32, device=dev)
        self.slot = torch.empty(cap, dtype=torch.int32, device=dev)
        self.delta_send = torch.empty(cap, dtype=table.dtype, device=dev)
        self.hist = torch.zeros(_lib.MAX_RANKS, dtype=torch.int64, device=dev)
        self.timeline = torch.zeros(8, dtype=torch.int64, device=dev)
        b = _lib.Bucket()
        for r in range(R):
            b.base[r] = self.buf.ptrs[r]
        b.off_inbox_keys, b.off_inbox_delta, b.off_inbox_meta, b.off_reply = off_keys, off_delta, off_meta, off_reply
        b.cap = cap
        b.send_keys = self.send_keys.data_ptr()
        b.slot = self.slot.data_ptr()
        b.delta_send = self.delta_send.data_ptr()
        b.hist = self.hist.data_ptr()
        b.timeline = self.timeline.data_ptr() if os.environ.get("MINIPS_BUCKET_TIMELINE") == "1" else None
        self.c = b
        self.launch_seq = 0
        self.last_epoch = -1        # table.write_epoch right after the last bucketed step

    def close(self):
        self.buf.close()


class SparseLRTrainer:
    """Device trainer bound to one worker's KVClientTable (fp32 or fp64 tables; the reference's
    tables are Val = double, lr_example.cpp:182)."""

    def __init__(self, client: KVClientTable, data: CsrBatchSource, alpha: float = 0.1):
        self.client = client
        self.table: DeviceTable = client.table
        self.data = data
        self.alpha = float(alpha)
        self.lib = client.lib
        self.dev = client.dev
        self._replica: Optional[torch.Tensor] = None
        self._gacc: Optional[torch.Tensor] = None
        self.avg_nnz = max(1, int(round(data.nnz / max(1, data.n_rows))))
        self._max_batch_nnz = -1
        self._bucket_batch = 0
        assert self.table.dtype in (torch.float32, torch.float64) and self.table.k == 1

    def _scratch(self):
        if self._replica is None:
            n = self.table.num_keys
            self._replica = torch.empty(n, dtype=self.table.dtype, device=self.dev)
            self._gacc = torch.zeros(n, dtype=self.table.dtype, device=self.dev)
        return self._replica, self._gacc

    def dense_regime(self, batch: int) -> bool:
        """Slab pull/push pays off once the batch touches a good part of the table."""
        nnz_per_row = max(1, self.data.nnz // max(1, self.data.n_rows))
        return batch * nnz_per_row * 4 >= self.table.num_keys

    def xchg_ok(self) -> bool:
        """The one-exchange-per-step kernel needs BSP and exactly one worker on every rank of the
        table, all training through it in lockstep."""
        t = self.table
        return (t.model_type == ModelType.BSP and t.num_workers == t.world and len(t.local_workers) == 1
                and t.hview is None)

    def bucket_ok(self) -> bool:
        """The message-flow step needs every rank of the table to run exactly one worker through it, step by
        step (any consistency model: it is BSP-tight, which satisfies the looser ones).  On one GPU every
        access is local already and the per-non-zero kernel is the right one."""
        t = self.table
        return (t.world > 1 and t.num_workers == t.world and len(t.local_workers) == 1 and t.hview is None
                and t.num_shards % t.world == 0)

    def _bucket(self) -> BucketScratch:
        t = self.table
        if getattr(t, "bucket", None) is None:
            # collective: every rank reaches its first bucketed step together; all agree on the capacity
            cap = max(t.control.all_gather(self._max_batch_nnz))
            t.bucket = BucketScratch(t, cap)
        return t.bucket

    def _xchg(self) -> XchgScratch:
        t = self.table
        if getattr(t, "xchg", None) is None:
            t.xchg = XchgScratch(t)        # collective: every rank reaches its first fused launch together
        return t.xchg

    @traced("lr.train_dense")
    def train_dense(self, batch_starts: torch.Tensor, batch: int, max_ctas: int = 0,
                    csr: Optional[_lib.Csr] = None, loss: Optional[torch.Tensor] = None,
                    timeline: Optional[torch.Tensor] = None, protocol: str = "auto") -> torch.Tensor:
        """Run len(batch_starts) iterations inside one persistent kernel.  Returns the per-
        iteration (logloss sum, #correct) tensor [n_iters, 2] (device, async).

        protocol: "xchg" = reduce at the owner, owner broadcasts (BSP, one worker per rank);
        "pushpull" = slab pull / slab push per iteration (any consistency model); "auto" picks
        xchg whenever it applies (every rank then has to make the same sequence of calls with the same
        number of iterations each: the kernel counts its hand-overs per launch).  Between two xchg launches
        on a table the OTHER ranks must not
        Add through the generic path (this rank's own generic writes are noticed and the replica
        is refreshed)."""
        c = self.client
        t = self.table
        if len(t.local_workers) > 1:
            raise RuntimeError("dense LR training is a full-occupancy cooperative kernel: one dense-training "
                               "worker per GPU and table (use step_sparse with several local workers)")
        n_iters = int(batch_starts.numel())
        if loss is None:
            loss = torch.zeros(n_iters, 2, dtype=torch.float32, device=self.dev)
        cv = t.clock_view()
        if csr is None:
            csr = self.data.c_struct()
        c._local_gate()
        tl_ptr = C.c_void_p(timeline.data_ptr()) if timeline is not None else None
        use_xchg = protocol == "xchg" or (protocol == "auto" and self.xchg_ok())
        if use_xchg:
            assert self.xchg_ok(), "xchg protocol: BSP table with exactly one worker on every rank"
            x = self._xchg()
            x.c.replica_valid = int(x.valid_clock == c.progress and x.valid_epoch == t.write_epoch)
            x.c.seq = x.launch_seq
            x.launch_seq += 1
            _lib.check(self.lib.mps_lr_train_xchg(
                C.byref(t.view), C.byref(cv), c.worker_id, c.progress, C.byref(csr),
                C.c_void_p(batch_starts.data_ptr()), n_iters, batch, self.alpha, C.byref(x.c),
                C.c_void_p(loss.data_ptr()), tl_ptr, self.avg_nnz, max_ctas, c._sp()), "lr_train_xchg")
            x.valid_clock = c.progress + n_iters
            x.valid_epoch = t.write_epoch
        else:
            replica, gacc = self._scratch()
            _lib.check(self.lib.mps_lr_train_dense_ex(
                C.byref(t.view), C.byref(cv), c.worker_id, c.progress, C.byref(csr),
                C.c_void_p(batch_starts.data_ptr()), n_iters, batch, self.alpha,
                C.c_void_p(replica.data_ptr()), C.c_void_p(gacc.data_ptr()), C.c_void_p(loss.data_ptr()),
                tl_ptr, self.avg_nnz, max_ctas, c._sp()), "lr_train_dense")
            t.write_epoch += 1      # in-place adds: a replica kept by the xchg kernel is stale now
        # host mirror of the clock (the kernel published the device-side one)
        with t.lock:
            c.progress += n_iters
            t.local_clock[c.worker_id] = c.progress
            ev = torch.cuda.Event()
            ev.record(c.stream or torch.cuda.current_stream())
            t.local_events[c.worker_id] = ev
            t.fold_clock = c.progress
            t.lock.notify_all()
        return loss

    @traced("lr.train_stream")
    def train_stream(self, loader, row_starts, host_loss: Optional[torch.Tensor] = None) -> torch.Tensor:
        """End-to-end iterations: every step copies its batch from pinned host memory (double
        buffered on a copy stream), trains on it, and returns its (logloss, #correct) to a
        pinned host tensor.  Returns that [n_steps, 2] host tensor (valid after a sync)."""
        n = len(row_starts)
        batch = loader.batch
        if host_loss is None:
            host_loss = torch.zeros(n, 2, dtype=torch.float32).pin_memory()
        dev_loss = torch.zeros(n, 2, dtype=torch.float32, device=self.dev)
        zero_start = torch.zeros(1, dtype=torch.int64, device=self.dev)
        stream = self.client.stream or torch.cuda.current_stream()
        pending = loader.prefetch(int(row_starts[0]))
        for i in range(n):
            k = pending
            if i + 1 < n:
                pending = loader.prefetch(int(row_starts[i + 1]))
            loader.acquire(k, stream)
            self.train_dense(zero_start, batch, csr=loader.staged_struct(k), loss=dev_loss[i:i + 1])
            loader.release(k, stream)
            host_loss[i].copy_(dev_loss[i], non_blocking=True)
        return host_loss

    @traced("lr.step_sparse")
    def step_sparse(self, row_begin: int, batch: int, loss: Optional[torch.Tensor] = None,
                    protocol: str = "auto") -> torch.Tensor:
        """One iteration in the sparse regime, then Clock.  protocol "bucket": keys, values and deltas travel
        between the GPUs as per-owner messages (lr_bucket.cu); "pernnz": one peer ld / red per non-zero
        (lr.cu); "auto" = bucket whenever it applies (several GPUs, one worker per rank, lockstep)."""
        c = self.client
        t = self.table
        if loss is None:
            loss = torch.zeros(2, dtype=torch.float32, device=self.dev)
        cv = t.clock_view()
        csr = self.data.c_struct()
        c._local_gate()
        wraps = row_begin + batch > self.data.n_rows
        use_bucket = protocol == "bucket" or (protocol == "auto" and self.bucket_ok() and not wraps)
        if use_bucket:
            assert self.bucket_ok() and not wraps, "bucket protocol: several ranks, one worker each, no wrap-around"
            if self._max_batch_nnz < 0:
                rp = self.data.row_ptr
                self._max_batch_nnz = int((rp[batch:] - rp[:-batch]).max().item()) if self.data.n_rows > batch \
                    else self.data.nnz
                self._bucket_batch = batch
            assert batch <= self._bucket_batch, "bucket scratch was sized for a smaller batch"
            bk = self._bucket()
            bk.c.seq = bk.launch_seq
            bk.launch_seq += 1
            gated = int(bk.last_epoch != t.write_epoch)       # previous clock came from the generic path
            _lib.check(self.lib.mps_lr_step_bucketed(
                C.byref(t.view), C.byref(cv), gated, c.progress, C.byref(csr), row_begin, batch, self.alpha,
                C.byref(bk.c), C.c_void_p(loss.data_ptr()), self.avg_nnz, c._sp()), "lr_step_bucketed")
            t.write_epoch += 1
            bk.last_epoch = t.write_epoch
            with t.lock:                                      # host mirror of the clock the kernel published
                c.progress += 1
                t.local_clock[c.worker_id] = c.progress
                ev = torch.cuda.Event()
                ev.record(c.stream or torch.cuda.current_stream())
                t.local_events[c.worker_id] = ev
                t.fold_clock = c.progress
                t.lock.notify_all()
            return loss
        else:
            _lib.check(self.lib.mps_lr_step_sparse(
                C.byref(t.view), C.byref(cv), c.worker_id, c.progress, C.byref(csr), row_begin, batch,
                self.alpha, C.c_void_p(loss.data_ptr()), c._sp()), "lr_step_sparse")
        t.write_epoch += 1
        c.clock()
        return loss

    def evaluate(self, weights: Optional[torch.Tensor] = None, n: Optional[int] = None) -> Tuple[float, float]:
        """(mean logloss, accuracy) of the local rows (reference test_error, lr_example.cpp:61-87)."""
        if weights is None:
            weights = self.client.get_all()
        n = self.data.n_rows if n is None else n
        out = torch.zeros(2, dtype=torch.float32, device=self.dev)
        csr = self.data.c_struct()
        t = self.table
        assert weights.dtype == t.dtype and weights.numel() == t.num_keys
        _lib.check(self.lib.mps_lr_eval(C.c_void_p(weights.data_ptr()),
            _lib.F32 if t.dtype == torch.float32 else _lib.F64,
                                        t.ranges[0][0], t.num_keys, C.byref(csr), 0, n,
                                        C.c_void_p(out.data_ptr()), self.client._sp()), "lr_eval")
        o = out.tolist()
        return o[0] / n, o[1] / n


# ------------------------------------------------------------------ fp32 PyTorch oracle
def oracle_step(w: torch.Tensor, data: CsrBatchSource, row_begin: int, batch: int, alpha: float
                ) -> Tuple[torch.Tensor, float, float]:
    """Gradient of one worker's batch at weights w (textbook mini-batch SGD; the reference's
    stale-delta accumulation quirk, SURVEY 2.8-2, is deliberately not reproduced).
    Returns (delta, logloss_sum, n_correct)."""
    n = data.n_rows
    idx = (torch.arange(batch, device=w.device) + row_begin) % n
    rp = data.row_ptr
    starts, ends = rp[idx], rp[idx + 1]
    lens = ends - starts
    row_of = torch.repeat_interleave(torch.arange(batch, device=w.device), lens)
    offs = torch.arange(int(lens.sum()), device=w.device) - torch.repeat_interleave(
        torch.cumsum(lens, 0) - lens, lens)
    nz = torch.repeat_interleave(starts, lens) + offs
    cols = data.col[nz].to(torch.int64) & 0xFFFFFFFF
    vals = data.val[nz]
    z = torch.zeros(batch, dtype=w.dtype, device=w.device).index_add_(0, row_of, w[cols] * vals.to(w.dtype))
    y = data.label[idx].to(w.dtype)
    vals = vals.to(w.dtype)
    pred = torch.sigmoid(z)
    g = alpha * (y - pred)
    delta = torch.zeros_like(w).index_add_(0, cols, g[row_of] * vals)
    loss = (torch.clamp(z, min=0) + torch.log1p(torch.exp(-z.abs())) - y * z).sum().item()
    correct = ((pred > 0.5) == (y > 0.5)).float().sum().item()
    return delta, loss, correct
