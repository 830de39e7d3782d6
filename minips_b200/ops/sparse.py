"""Device versions of the reference's host-side sparse helpers (csrc/kernels/sparse_utils.cu):
the per-batch sorted unique key set (lib/batch_data_sampler.cpp:50-71) and libsvm text -> CSR
(lib/parser.hpp:17-47)."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib

_TILE = 4096


def batch_unique_keys(data, row_begin: int, batch: int, num_dims: int, want_slots: bool = False,
                      scratch: Optional[torch.Tensor] = None):
    """Sorted unique feature ids of rows [row_begin, row_begin+batch) of a CsrBatchSource.
    Returns (keys int32[count], slots int32[batch nnz] | None)."""
    lib = _lib.load()
    dev = data.col.device
    words = (num_dims + 31) // 32
    need = 2 * words + (words + _TILE - 1) // _TILE + 2
    if scratch is None or scratch.numel() < need:
        scratch = torch.empty(need, dtype=torch.int32, device=dev)
    rp = data.row_ptr
    nnz_lo = int(rp[row_begin]) if row_begin + batch <= data.n_rows else 0
    nnz_hi = int(rp[min(row_begin + batch, data.n_rows)])
    max_keys = min(num_dims, (nnz_hi - nnz_lo) if row_begin + batch <= data.n_rows else data.nnz)
    keys = torch.empty(max(1, max_keys), dtype=torch.int32, device=dev)
    count = torch.zeros(1, dtype=torch.int32, device=dev)
    slots = None
    if want_slots:
        assert row_begin + batch <= data.n_rows, "slot map needs a non-wrapping batch"
        slots = torch.empty(max(1, nnz_hi - nnz_lo), dtype=torch.int32, device=dev)
    csr = data.c_struct()
    _lib.check(lib.mps_batch_unique_keys(
        C.byref(csr), row_begin, batch, num_dims, C.c_void_p(scratch.data_ptr()), C.c_void_p(keys.data_ptr()),
        C.c_void_p(count.data_ptr()), C.c_void_p(slots.data_ptr()) if slots is not None else None, nnz_lo,
        _lib.stream_ptr()), "batch_unique_keys")
    n = int(count.item())
    return keys[:n], slots


def libsvm_to_csr(text: torch.Tensor, map_negative_label_to_zero: bool = True):
    """Parse a uint8 CUDA tensor holding libsvm text into a CsrBatchSource, entirely on device."""
    from ..models.lr import CsrBatchSource
    lib = _lib.load()
    assert text.is_cuda and text.dtype == torch.uint8
    dev = text.device
    nbytes = text.numel()
    sp = _lib.stream_ptr()
    counts = torch.zeros(4, dtype=torch.int32, device=dev)
    scratch = torch.empty((nbytes + 1 + _TILE - 1) // _TILE + 4, dtype=torch.int32, device=dev)
    # an upper bound on the number of lines without a pass of its own: count '\n' with torch
    max_lines = int((text == 10).sum().item()) + 1
    line_end = torch.empty(max_lines, dtype=torch.int64, device=dev)
    tp = C.c_void_p(text.data_ptr())
    _lib.check(lib.mps_libsvm_scan_lines(tp, nbytes, C.c_void_p(line_end.data_ptr()), C.c_void_p(scratch.data_ptr()),
                                         C.c_void_p(counts.data_ptr()), sp), "libsvm_scan_lines")
    n_lines = int(counts[0].item())
    valid = torch.empty(n_lines, dtype=torch.int32, device=dev)
    nnz = torch.empty(n_lines, dtype=torch.int32, device=dev)
    row_of_line = torch.empty(n_lines, dtype=torch.int64, device=dev)
    nnz_prefix = torch.empty(n_lines, dtype=torch.int64, device=dev)
    row_ptr = torch.zeros(n_lines + 1, dtype=torch.int64, device=dev)
    scratch2 = torch.empty((n_lines + _TILE - 1) // _TILE + 4, dtype=torch.int32, device=dev)
    _lib.check(lib.mps_libsvm_index_lines(
        tp, C.c_void_p(line_end.data_ptr()), n_lines, C.c_void_p(valid.data_ptr()), C.c_void_p(nnz.data_ptr()),
        C.c_void_p(row_of_line.data_ptr()), C.c_void_p(nnz_prefix.data_ptr()), C.c_void_p(row_ptr.data_ptr()),
        C.c_void_p(scratch2.data_ptr()), C.c_void_p(counts.data_ptr()), sp), "libsvm_index_lines")
    _, rows, total = [int(v) for v in counts[:3].tolist()]
    col = torch.empty(max(1, total), dtype=torch.int32, device=dev)
    val = torch.empty(max(1, total), dtype=torch.float32, device=dev)
    label = torch.empty(max(1, rows), dtype=torch.float32, device=dev)
    _lib.check(lib.mps_libsvm_parse(
        tp, C.c_void_p(line_end.data_ptr()), n_lines, C.c_void_p(row_of_line.data_ptr()),
        C.c_void_p(nnz_prefix.data_ptr()), C.c_void_p(col.data_ptr()), C.c_void_p(val.data_ptr()),
        C.c_void_p(label.data_ptr()), int(map_negative_label_to_zero), sp), "libsvm_parse")
    row_ptr = row_ptr[:rows + 1].clone()
    row_ptr[rows] = total
    return CsrBatchSource(row_ptr, col[:total], val[:total], label[:rows])


def load_libsvm_file(path: str, device="cuda", map_negative_label_to_zero: bool = True):
    """Local libsvm file -> CSR in HBM (the HDFS loader's replacement for real files)."""
    import numpy as np
    raw = np.fromfile(path, dtype=np.uint8)
    return libsvm_to_csr(torch.from_numpy(raw).to(device), map_negative_label_to_zero)


def write_synthetic_libsvm(path: str, n_rows: int, n_features: int, nnz_per_row: int = 10, seed: int = 0):
    """Small synthetic libsvm text file (BASELINE.json config 1: 10k samples x 1k features)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    w = rng.standard_normal(n_features)
    with open(path, "w") as f:
        for _ in range(n_rows):
            idx = np.sort(rng.choice(n_features, size=nnz_per_row, replace=False))
            v = rng.standard_normal(nnz_per_row)
            y = 1 if (w[idx] * v).sum() + 0.3 * rng.standard_normal() > 0 else -1
            f.write(("+1" if y > 0 else "-1") + " " + " ".join(f"{i + 1}:{x:.6g}" for i, x in zip(idx, v)) + "\n")
    return path
