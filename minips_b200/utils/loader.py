"""Double-buffered host->device batch loader.

The reference only sketches this component (lib/abstract_aync_data_loader.hpp:35-92 is an
empty skeleton of an AsyncReadBuffer that prefetches batches on a side thread).  Here the
"side thread" is a CUDA copy stream: batch i+1 is copied from pinned host memory while the
training kernel of batch i runs; an event hands the staged buffers to the compute stream.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from ..models.lr import CsrBatchSource


class PackedCsrHost:
    """Compact pinned-host form of a CSR data set for the H2D stream: row lengths as u8, feature ids
    bit-packed at 16 / 20 / 24 bits (the smallest that holds every id; little-endian bit stream), labels as
    u8, values unchanged (fp32).  Lossless when every row has < 256 non-zeros, ids fit 24 bits and labels
    are 0/1; `from_csr` returns None otherwise."""

    def __init__(self, row_ptr, row_len, ids, id_bits, val, label_u8):
        self.row_ptr, self.row_len, self.ids, self.id_bits = row_ptr, row_len, ids, id_bits
        self.val, self.label_u8 = val, label_u8
        self.n_rows = row_len.numel()
        self.nnz = val.numel()

    @staticmethod
    def pack_ids(col: torch.Tensor, id_bits: int) -> torch.Tensor:
        """int32 ids -> uint8 little-endian bit stream of id_bits per id (plus 8 bytes of slack)."""
        col = col.to(torch.int32).contiguous()
        if id_bits == 16:
            body = col.to(torch.int16).view(torch.uint8)
        elif id_bits == 24:
            body = col.view(torch.uint8).view(-1, 4)[:, :3].contiguous().view(-1)
        elif id_bits == 20:
            if col.numel() % 2:
                col = torch.cat([col, col.new_zeros(1)])
            pairs = col.view(-1, 2).to(torch.int64)
            body = (pairs[:, 0] | (pairs[:, 1] << 20)).view(torch.uint8).view(-1, 8)[:, :5].contiguous().view(-1)
        else:
            raise ValueError(id_bits)
        return torch.cat([body, body.new_zeros(8)])

    @staticmethod
    def from_csr(host: CsrBatchSource) -> "Optional[PackedCsrHost]":
        rp = host.row_ptr.cpu()
        lens = rp[1:] - rp[:-1]
        col = host.col.cpu()
        lab = host.label.cpu()
        cmax = int(col.max()) if col.numel() else 0
        if int(lens.max()) > 255 or cmax >= (1 << 24) or int(col.min()) < 0:
            return None
        if not bool(((lab == 0) | (lab == 1)).all()):
            return None
        id_bits = 16 if cmax < (1 << 16) else (20 if cmax < (1 << 20) else 24)
        return PackedCsrHost(rp.contiguous(), lens.to(torch.uint8).pin_memory(),
                             PackedCsrHost.pack_ids(col, id_bits).pin_memory(), id_bits,
                             host.val.cpu().contiguous().pin_memory(), lab.to(torch.uint8).pin_memory())


class AsyncBatchLoader:
    """depth staged batches on the device, filled by a copy stream from pinned host memory.

    in_place=True (packed host data only): the training kernel reads the compact form directly
    (csrc/kernels/lr_common.cuh CsrView<true>), so a step needs no device-side rebuild — the loader
    ships, besides the packed arrays, the nnz offset of every 8th row of the batch (64 KiB), which it
    cuts out of the host row_ptr."""

    def __init__(self, host, batch: int, device, depth: int = 2, in_place: bool = False):
        self.packed = isinstance(host, PackedCsrHost)
        if self.packed:
            assert host.row_len.is_pinned() and host.ids.is_pinned() and host.val.is_pinned()
            assert host.label_u8.is_pinned()
        else:
            assert host.row_ptr.is_pinned() and host.col.is_pinned() and host.val.is_pinned() and host.label.is_pinned()
        self.host = host
        self.in_place = bool(in_place) and self.packed
        assert not self.in_place or batch % 8 == 0
        self.batch = batch
        self.device = device
        self.depth = depth
        # one copy stream; MINIPS_COPY_STREAMS=2 alternates two per batch (measured on B200: no difference,
        # 334.6 against 333.9 us per 17.2 MB step - the link is the limit, not the gaps between copies)
        n_cs = max(1, int(os.environ.get("MINIPS_COPY_STREAMS", "1")))
        self.copy_streams = [torch.cuda.Stream(device) for _ in range(n_cs)]
        self.copy_stream = self.copy_streams[0]
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
            if self.packed:
                self.slots[-1].update({
                    "row_len": torch.empty(batch, dtype=torch.uint8, device=device),
                    "ids": torch.empty((max_nnz * host.id_bits + 7) // 8 + 16, dtype=torch.uint8, device=device),
                    "label_u8": torch.empty(batch, dtype=torch.uint8, device=device),
                    "scratch": torch.empty(batch // 4096 + 4, dtype=torch.int32, device=device),
                })
                if self.in_place:
                    # the three small arrays of a batch (coarse offsets, row lengths, labels) travel as ONE copy:
                    # a pinned staging area per slot, laid out like the device buffer they land in
                    nc = 8 * (batch // 8 + 1)
                    small_dev = torch.empty(nc + 2 * batch, dtype=torch.uint8, device=device)
                    small_host = torch.empty(nc + 2 * batch, dtype=torch.uint8).pin_memory()
                    self.slots[-1].update({
                        "small": small_dev, "small_host": small_host,
                        "coarse": small_dev[:nc].view(torch.int64), "coarse_host": small_host[:nc].view(torch.int64),
                        "row_len": small_dev[nc:nc + batch], "row_len_host": small_host[nc:nc + batch],
                        "label_u8": small_dev[nc + batch:], "label_u8_host": small_host[nc + batch:],
                        "copied": None,
                    })
        self.bytes_per_batch = 0
        # device-side CSR rebuild (scan x3 + unpack) unless the kernel reads the packed form itself
        self.launches_per_batch = 4 if (self.packed and not self.in_place) else 0
        self._i = 0

    def prefetch(self, row_start: int) -> int:
        """Enqueue the H2D copy of rows [row_start, row_start + batch) into the next slot."""
        h, B = self.host, self.batch
        assert row_start + B <= h.n_rows
        k = self._i % self.depth
        cs = self.copy_streams[self._i % len(self.copy_streams)]
        self._i += 1
        slot = self.slots[k]
        b, e = int(h.row_ptr[row_start]), int(h.row_ptr[row_start + B])
        with torch.cuda.stream(cs):
            if slot["free"] is not None:
                cs.wait_event(slot["free"])
            if self.packed:
                if not self.in_place:
                    slot["row_len"].copy_(h.row_len[row_start:row_start + B], non_blocking=True)
                w = h.id_bits
                byte0, byte1 = (b * w) // 8, (e * w + 7) // 8 + 4   # + the word of slack the unpack kernel may touch
                slot["ids"][:byte1 - byte0].copy_(h.ids[byte0:byte1], non_blocking=True)
                slot["bit_off"] = (b * w) % 8
                slot["id_bytes"] = byte1 - byte0
                slot["val"][:e - b].copy_(h.val[b:e], non_blocking=True)
                if self.in_place:
                    if slot["copied"] is not None:
                        slot["copied"].synchronize()      # the previous copy out of this pinned staging area
                    torch.sub(h.row_ptr[row_start:row_start + B + 1:8], b, out=slot["coarse_host"])
                    slot["row_len_host"].copy_(h.row_len[row_start:row_start + B])
                    slot["label_u8_host"].copy_(h.label_u8[row_start:row_start + B])
                    slot["small"].copy_(slot["small_host"], non_blocking=True)
                    slot["copied"] = torch.cuda.Event()
                    slot["copied"].record(cs)
                else:
                    slot["label_u8"].copy_(h.label_u8[row_start:row_start + B], non_blocking=True)
            else:
                slot["row_ptr"].copy_(h.row_ptr[row_start:row_start + B + 1], non_blocking=True)
                slot["col"][:e - b].copy_(h.col[b:e], non_blocking=True)
                slot["val"][:e - b].copy_(h.val[b:e], non_blocking=True)
                slot["label"].copy_(h.label[row_start:row_start + B], non_blocking=True)
            slot["ready"].record(cs)
        slot["base"] = 0 if self.packed else b
        slot["nnz"] = e - b
        self.bytes_per_batch = (B + slot["id_bytes"] + 4 * (e - b) + B + (8 * (B // 8 + 1) if self.in_place else 0)) \
            if self.packed else ((B + 1) * 8 + (e - b) * 8 + B * 4)
        return k

    def acquire(self, k: int, stream: Optional[torch.cuda.Stream] = None) -> CsrBatchSource:
        """Make `stream` wait for slot k and return a CSR view of it.  The staged row_ptr keeps
        the ABSOLUTE nnz offsets of the host arrays, so col/val are exposed through pointers
        shifted back by the slot's first offset (see staged_struct)."""
        slot = self.slots[k]
        st = stream or torch.cuda.current_stream()
        st.wait_event(slot["ready"])
        if self.packed and not self.in_place:
            # rebuild row_ptr / col / label on the compute stream (three small scan launches + one unpack)
            import ctypes as C
            from ..ops import _lib
            lib = _lib.load()
            _lib.check(lib.mps_csr_unpack(
                C.c_void_p(slot["row_len"].data_ptr()), C.c_void_p(slot["ids"].data_ptr()), self.host.id_bits,
                slot["bit_off"], C.c_void_p(slot["label_u8"].data_ptr()), self.batch, slot["nnz"],
                C.c_void_p(slot["row_ptr"].data_ptr()), C.c_void_p(slot["col"].data_ptr()),
                C.c_void_p(slot["label"].data_ptr()), C.c_void_p(slot["scratch"].data_ptr()),
                _lib.stream_ptr(st)), "csr_unpack")
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
        if self.in_place:
            m.format = _lib.CSR_PACKED
            m.id_bits = self.host.id_bits
            m.bit_off = slot["bit_off"]
            m.row_len = slot["row_len"].data_ptr()
            m.coarse = slot["coarse"].data_ptr()
            m.packed_ids = slot["ids"].data_ptr()
            m.label_u8 = slot["label_u8"].data_ptr()
        return m

    def release(self, k: int, stream: Optional[torch.cuda.Stream] = None):
        ev = torch.cuda.Event()
        ev.record(stream or torch.cuda.current_stream())
        self.slots[k]["free"] = ev
