"""Double-buffered host->device batch loader.

The reference only sketches this component (lib/abstract_aync_data_loader.hpp:35-92 is an
empty skeleton of an AsyncReadBuffer that prefetches batches on a side thread).  Here the
"side thread" is a CUDA copy stream: batch i+1 is copied from pinned host memory while the
training kernel of batch i runs; an event hands the staged buffers to the compute stream.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ..models.lr import CsrBatchSource


class AsyncBatchLoader:
    def __init__(self, host: CsrBatchSource, batch: int, device, depth: int = 2):
        assert host.row_ptr.is_pinned() and host.col.is_pinned() and host.val.is_pinned() and host.label.is_pinned()
        self.host = host
        self.batch = batch
        self.device = device
        self.depth = depth
        self.copy_stream = torch.cuda.Stream(device)
        # worst-case nnz of any `batch` consecutive rows
        rp = host.row_ptr
        max_nnz = int((rp[batch:] - rp[:-batch]).max().item()) if host.n_rows > batch else host.nnz
        self.slots = []
        for _ in range(depth):
            self.slots.append({
                "row_ptr": torch.empty(batch + 1, dtype=torch.int64, device=device),
                "col": torch.empty(max_nnz, dtype=torch.int32, device=device),
                "val": torch.empty(max_nnz, dtype=torch.float32, device=device),
                "label": torch.empty(batch, dtype=torch.float32, device=device),
                "ready": torch.cuda.Event(),
                "free": None,
            })
        self.bytes_per_batch = 0
        self._i = 0

    def prefetch(self, row_start: int) -> int:
        """Enqueue the H2D copy of rows [row_start, row_start + batch) into the next slot."""
        h, B = self.host, self.batch
        assert row_start + B <= h.n_rows
        k = self._i % self.depth
        self._i += 1
        slot = self.slots[k]
        b, e = int(h.row_ptr[row_start]), int(h.row_ptr[row_start + B])
        with torch.cuda.stream(self.copy_stream):
            if slot["free"] is not None:
                self.copy_stream.wait_event(slot["free"])
            slot["row_ptr"].copy_(h.row_ptr[row_start:row_start + B + 1], non_blocking=True)
            slot["col"][:e - b].copy_(h.col[b:e], non_blocking=True)
            slot["val"][:e - b].copy_(h.val[b:e], non_blocking=True)
            slot["label"].copy_(h.label[row_start:row_start + B], non_blocking=True)
            slot["ready"].record(self.copy_stream)
        slot["base"] = b
        slot["nnz"] = e - b
        self.bytes_per_batch = (B + 1) * 8 + (e - b) * 8 + B * 4
        return k

    def acquire(self, k: int, stream: Optional[torch.cuda.Stream] = None) -> CsrBatchSource:
        """Make `stream` wait for slot k and return a CSR view of it.  The staged row_ptr keeps
        the ABSOLUTE nnz offsets of the host arrays, so col/val are exposed through pointers
        shifted back by the slot's first offset (see staged_struct)."""
        slot = self.slots[k]
        (stream or torch.cuda.current_stream()).wait_event(slot["ready"])
        return CsrBatchSource(slot["row_ptr"], slot["col"], slot["val"], slot["label"])

    def staged_struct(self, k: int):
        from ..ops import _lib
        slot = self.slots[k]
        m = _lib.Csr()
        m.row_ptr = slot["row_ptr"].data_ptr()
        m.col = slot["col"].data_ptr() - 4 * slot["base"]
        m.val = slot["val"].data_ptr() - 4 * slot["base"]
        m.label = slot["label"].data_ptr()
        m.n_rows = self.batch
        return m

    def release(self, k: int, stream: Optional[torch.cuda.Stream] = None):
        ev = torch.cuda.Event()
        ev.record(stream or torch.cuda.current_stream())
        self.slots[k]["free"] = ev
