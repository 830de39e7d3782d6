"""ctypes binding of libminips_kernels.so (C ABI in include/minips_kernels.h).

The library is built in-tree by build.py (nvcc, sm_100a).  Importing this module never
needs a GPU; calling an op does.  If a CUDA device is present and the library is missing,
loading fails loudly instead of falling back to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

MAX_SHARDS = 64
MAX_RANKS = 16
MAX_WORKERS = 256
PAD_PROGRESS = 0
PAD_PULLED = PAD_PROGRESS + MAX_WORKERS
PAD_FOLDED = PAD_PULLED + MAX_WORKERS
PAD_BARRIER = PAD_FOLDED + MAX_SHARDS
PAD_ERROR = PAD_BARRIER + MAX_RANKS
PAD_MINCLOCK = PAD_ERROR + 1
PAD_LOCALCTR = PAD_MINCLOCK + 1
PAD_XCNT = PAD_LOCALCTR + 4
PAD_INTS = 1024

MODEL_SSP, MODEL_BSP, MODEL_ASP = 0, 1, 2
F32, F64 = 0, 1
CSR_PLAIN, CSR_PACKED = 0, 1
XCHG_REDUCE_P2P = 1
XCHG_BCAST_P2P = 2


class TableView(C.Structure):
    _fields_ = [
        ("weights", C.c_void_p * MAX_SHARDS),
        ("staging", C.c_void_p * MAX_SHARDS),
        ("begin", C.c_uint32 * (MAX_SHARDS + 1)),
        ("num_shards", C.c_int32),
        ("uniform_width", C.c_uint32),
        ("vals_per_key", C.c_int32),
        ("dtype", C.c_int32),
    ]


class ClockView(C.Structure):
    _fields_ = [
        ("pad", C.c_void_p * MAX_RANKS),
        ("num_ranks", C.c_int32),
        ("my_rank", C.c_int32),
        ("num_workers", C.c_int32),
        ("num_shards", C.c_int32),
        ("own_first", C.c_int32),
        ("own_count", C.c_int32),
        ("model", C.c_int32),
        ("staleness", C.c_int32),
        ("timeout_cycles", C.c_int64),
    ]


class Csr(C.Structure):
    _fields_ = [
        ("row_ptr", C.c_void_p),
        ("col", C.c_void_p),
        ("val", C.c_void_p),
        ("label", C.c_void_p),
        ("n_rows", C.c_int64),
        ("format", C.c_int32),      # CSR_PLAIN / CSR_PACKED
        ("id_bits", C.c_int32),
        ("bit_off", C.c_int32),
        ("reserved", C.c_int32),
        ("row_len", C.c_void_p),
        ("coarse", C.c_void_p),
        ("packed_ids", C.c_void_p),
        ("label_u8", C.c_void_p),
    ]


class Xchg(C.Structure):
    """MpsXchg: symmetric scratch of the reduce-at-owner / owner-broadcasts LR kernel."""
    _fields_ = [
        ("replica", C.c_void_p * MAX_RANKS),
        ("gacc", C.c_void_p * MAX_RANKS),
        ("replica_mc", C.c_void_p),
        ("gacc_mc", C.c_void_p),
        ("gacc_stride", C.c_int64),
        ("replica_valid", C.c_int32),
        ("flags", C.c_int32),
        ("seq", C.c_int32),
        ("reserved", C.c_int32),
    ]


class Bucket(C.Structure):
    """MpsBucket: scratch of the bucketed (message-flow) sparse LR step."""
    _fields_ = [
        ("base", C.c_void_p * MAX_RANKS),
        ("off_inbox_keys", C.c_int64),
        ("off_inbox_delta", C.c_int64),
        ("off_inbox_meta", C.c_int64),
        ("off_reply", C.c_int64),
        ("cap", C.c_int64),
        ("send_keys", C.c_void_p),
        ("slot", C.c_void_p),
        ("delta_send", C.c_void_p),
        ("hist", C.c_void_p),
        ("timeline", C.c_void_p),
        ("seq", C.c_int32),
        ("reserved", C.c_int32),
    ]


class HashView(C.Structure):
    _fields_ = [
        ("keys", C.c_void_p * MAX_SHARDS),
        ("vals", C.c_void_p * MAX_SHARDS),
        ("capacity", C.c_uint32 * MAX_SHARDS),
        ("begin", C.c_uint32 * (MAX_SHARDS + 1)),
        ("num_shards", C.c_int32),
    ]


_LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libminips_kernels.so"
_lib = None


class KernelLibraryMissing(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def load(build_if_missing: bool = True):
    """Load (building first if needed) the kernel library."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists() and build_if_missing and os.environ.get("MINIPS_NO_AUTOBUILD") != "1":
        import importlib.util
        spec = importlib.util.spec_from_file_location("minips_build", _LIB_PATH.parents[2] / "build.py")
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build_kernels(verbose=False)
    if not _LIB_PATH.exists():
        raise KernelLibraryMissing(
            f"{_LIB_PATH} not built; run `python build.py` (there is no PyTorch fallback for device ops)")
    lib = C.CDLL(str(_LIB_PATH))
    _declare(lib)
    _lib = lib
    return lib


def _declare(lib):
    vp, i32, i64, f32, f64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double
    P = C.POINTER
    sig = {
        "mps_device_count": ([], i32),
        "mps_set_device": ([i32], i32),
        "mps_sm_count": ([i32], i32),
        "mps_heap_alloc": ([C.c_size_t, P(vp)], i32),
        "mps_heap_free": ([vp], i32),
        "mps_ipc_get_handle": ([vp, C.c_char_p], i32),
        "mps_ipc_open_handle": ([C.c_char_p, P(vp)], i32),
        "mps_ipc_close_handle": ([vp], i32),
        "mps_enable_peer": ([i32], i32),
        "mps_memset32_async": ([vp, i32, C.c_size_t, vp], i32),
        "mps_error_string": ([i32], C.c_char_p),
        "mps_pull": ([P(TableView), P(ClockView), i32, i32, vp, vp, i64, vp], i32),
        "mps_push": ([P(TableView), vp, vp, i64, vp], i32),
        "mps_pull_dense": ([P(TableView), P(ClockView), i32, i32, vp, vp], i32),
        "mps_push_dense": ([P(TableView), vp, i32, i32, vp], i32),
        "mps_clock": ([P(TableView), P(ClockView), i32, i32, i32, vp], i32),
        "mps_fold_server": ([P(TableView), P(ClockView), i32, vp, vp], i32),
        "mps_barrier": ([P(ClockView), i32, vp], i32),
        "mps_reset_pad": ([P(ClockView), i32, vp], i32),
        "mps_straggle": ([C.c_longlong, vp], i32),
        "mps_lr_step_sparse": ([P(TableView), P(ClockView), i32, i32, P(Csr), i64, i64, f64, vp, vp], i32),
        "mps_lr_step_bucketed": ([P(TableView), P(ClockView), i32, i32, P(Csr), i64, i64, f64, P(Bucket), vp, i32,
                                  vp], i32),
        "mps_lr_train_dense": ([P(TableView), P(ClockView), i32, i32, P(Csr), vp, i32, i64, f64, vp, vp, vp,
                                i32, vp], i32),
        "mps_lr_train_dense_ex": ([P(TableView), P(ClockView), i32, i32, P(Csr), vp, i32, i64, f64, vp, vp, vp,
                                   vp, i32, i32, vp], i32),
        "mps_lr_train_xchg": ([P(TableView), P(ClockView), i32, i32, P(Csr), vp, i32, i64, f64, P(Xchg), vp, vp,
                               i32, i32, vp], i32),
        "mps_lr_eval": ([vp, i32, C.c_uint32, C.c_uint32, P(Csr), i64, i64, vp, vp], i32),
        "mps_batch_unique_keys": ([P(Csr), i64, i64, C.c_uint32, vp, vp, vp, vp, i64, vp], i32),
        "mps_libsvm_scan_lines": ([vp, i64, vp, vp, vp, vp], i32),
        "mps_libsvm_index_lines": ([vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp], i32),
        "mps_libsvm_parse": ([vp, vp, i64, vp, vp, vp, vp, vp, i32, vp], i32),
        "mps_csr_unpack": ([vp, vp, i32, i32, vp, i64, i64, vp, vp, vp, vp, vp], i32),
        "mps_hash_clear": ([vp, vp, i64, vp], i32),
        "mps_hash_add": ([P(HashView), vp, vp, i64, vp, vp], i32),
        "mps_hash_get": ([P(HashView), P(ClockView), i32, vp, vp, i64, vp], i32),
        "mps_hash_export": ([vp, vp, i64, vp, vp, vp, vp], i32),
        "mps_kmeans_assign_tc": ([P(TableView), P(ClockView), i32, vp, vp, i64, i32, i32, i32, i32, vp, vp, vp, vp, vp,
                                  vp, i32, vp], i32),
        "mps_dense_lr_forward_tc": ([P(TableView), P(ClockView), i32, vp, vp, vp, i64, i32, f32, vp, vp, vp, vp, vp,
                                     i32, vp], i32),
        "mps_dense_lr_step_tc": ([P(TableView), P(ClockView), i32, vp, vp, vp, i64, i32, f32, vp, vp, vp, vp, vp,
                                  i32, vp], i32),
        "mps_dense_lr_step_stream": ([P(TableView), P(ClockView), i32, vp, vp, vp, i64, i32, f32, vp, vp, vp, i32, vp],
                                     i32),
        "mps_dense_lr_step_rows": ([P(TableView), P(ClockView), i32, vp, vp, vp, i64, i32, f32, vp, vp, vp, i32, vp],
                                   i32),
        "mps_dense_lr_grad": ([vp, vp, i64, i32, vp, vp], i32),
        "mps_kmeans_tc_error": ([], i32),
        "mps_dense_lr_tc_error": ([], i32),
        "mps_kmeans_assign_ref": ([vp, vp, vp, i64, i32, i32, vp, vp, vp], i32),
        "mps_kmeans_accumulate": ([vp, vp, vp, i64, i32, i32, vp, vp, vp, vp], i32),
        "mps_kmeans_delta": ([vp, vp, vp, vp, i32, i32, f32, vp, vp, vp], i32),
        "mps_kmeans_delta_push": ([P(TableView), P(TableView), P(ClockView), i32, vp, vp, vp, i32, i32, i32, i32, f32,
                                   vp], i32),
        "mps_kmeans_step": ([P(TableView), P(ClockView), P(TableView), P(ClockView), i32, i32, i32, vp, i64, i32, i32,
                             i32, i32, f32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp], i32),
        "mps_kmeans_accumulate_sorted": ([vp, vp, vp, i64, i32, i32, vp, vp, vp, vp, vp], i32),
    }
    for name, (argtypes, restype) in sig.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.argtypes = argtypes
            fn.restype = restype


class CudaError(RuntimeError):
    pass


def check(err: int, what: str = ""):
    if err != 0:
        msg = load().mps_error_string(err)
        raise CudaError(f"{what}: CUDA error {err}: {msg.decode() if msg else '?'}")


def stream_ptr(stream=None) -> C.c_void_p:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
