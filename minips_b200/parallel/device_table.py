"""Device-resident parameter-server tables.

A table is range-partitioned into shards (reference driver/engine.hpp:289-357: one storage +
one consistency model per server thread).  Here a shard is a region of its owner's HBM inside
a symmetric slab; "server threads" do not exist — peers read / accumulate into the shard
directly (csrc/kernels/push_pull.cu) and the consistency model is a set of flags in the
slab's signal pad (csrc/kernels/clock.cu).
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Dict, List, Optional, Sequence

import torch

from ..ops import _lib
from .heap import SymmetricBuffer, _ControlPlane
from .types import ModelType, Range, StorageType

_ALIGN = 256


def _align(x: int, a: int = _ALIGN) -> int:
    return (x + a - 1) // a * a


def keys_as_i32(keys: torch.Tensor, device) -> torch.Tensor:
    """uint32 key bit patterns carried in an int32 CUDA tensor (reference Key = uint32_t,
    base/magic.hpp:8)."""
    if not isinstance(keys, torch.Tensor):
        keys = torch.as_tensor(keys, dtype=torch.int64)
    if keys.dtype == torch.int32:
        k = keys
    elif keys.dtype == torch.uint32:
        k = keys.view(torch.int32)
    else:
        k = keys.to(torch.int64).bitwise_and(0xFFFFFFFF)
        k = torch.where(k >= 2 ** 31, k - 2 ** 32, k).to(torch.int32)
    return k.to(device, non_blocking=True).contiguous()


class DeviceTable:
    def __init__(self, table_id: int, ranges: Sequence[Range], shard_owner: Sequence[int],
                 control: _ControlPlane, model_type: ModelType = ModelType.SSP, staleness: int = 0,
                 storage_type: StorageType = StorageType.Vector, dtype: torch.dtype = torch.float32,
                 vals_per_key: int = 1, device: Optional[int] = None, hash_capacity: Optional[int] = None):
        assert len(ranges) == len(shard_owner) and 0 < len(ranges) <= _lib.MAX_SHARDS
        assert dtype in (torch.float32, torch.float64)
        self.lib = _lib.load()
        self.table_id = table_id
        self.ranges = [(int(b), int(e)) for b, e in ranges]
        for (b0, e0), (b1, _) in zip(self.ranges, self.ranges[1:]):
            assert e0 == b1, "shard ranges must be contiguous and ascending"
        self.shard_owner = list(shard_owner)
        assert self.shard_owner == sorted(self.shard_owner), "shards are laid out rank-major"
        self.control = control
        self.rank, self.world = control.rank, control.world
        assert self.world <= _lib.MAX_RANKS
        self.model_type = ModelType(model_type)
        self.staleness = int(staleness)
        self.storage_type = StorageType(storage_type)
        self.dtype = dtype
        self.itemsize = 4 if dtype == torch.float32 else 8
        self.k = int(vals_per_key)
        self.device = torch.cuda.current_device() if device is None else device
        self.num_shards = len(self.ranges)
        self.num_keys = self.ranges[-1][1] - self.ranges[0][0]

        # identical slab layout on every rank (sized for the largest rank): [weights|staging] per
        # owned shard, then the signal pad
        self.woff: List[int] = [0] * self.num_shards
        self.soff: List[int] = [0] * self.num_shards
        per_rank = [0] * self.world
        bsp = self.model_type == ModelType.BSP
        # MapStorage on device (server/map_storage.hpp): every shard is an open-addressing hash of
        # `hash_capacity` slots instead of a dense array over its key range
        self.hash_capacity = None
        if hash_capacity:
            cap = 1
            while cap < hash_capacity:
                cap <<= 1
            self.hash_capacity = cap
            assert not bsp, "hash-backed (Map) tables add in place: use SSP or ASP"
            assert dtype == torch.float32 and self.k == 1
        for s, (b, e) in enumerate(self.ranges):
            r = self.shard_owner[s]
            nbytes = (_align((e - b) * self.k * self.itemsize) if not self.hash_capacity
                      else _align(self.hash_capacity * 8))
            self.woff[s] = per_rank[r]
            per_rank[r] += nbytes
            if bsp:
                self.soff[s] = per_rank[r]
                per_rank[r] += nbytes
        self.pad_off = _align(max(per_rank))
        self.slab = SymmetricBuffer(self.pad_off + _lib.PAD_INTS * 4, control, self.device)

        self.own_shards = [s for s in range(self.num_shards) if self.shard_owner[s] == self.rank]
        self.own_first = self.own_shards[0] if self.own_shards else 0
        self.own_count = len(self.own_shards)

        v = _lib.TableView()
        for s in range(self.num_shards):
            base = self.slab.ptrs[self.shard_owner[s]]
            v.weights[s] = base + self.woff[s]
            v.staging[s] = (base + self.soff[s]) if bsp else None
            v.begin[s] = self.ranges[s][0]
        v.begin[self.num_shards] = self.ranges[-1][1]
        v.num_shards = self.num_shards
        widths = [e - b for b, e in self.ranges]
        uw = widths[0]
        uniform = uw > 0 and all(w == uw for w in widths[:-1]) and widths[-1] >= uw
        v.uniform_width = uw if (uniform and self.num_shards > 1) else 0
        v.vals_per_key = self.k
        v.dtype = _lib.F32 if dtype == torch.float32 else _lib.F64
        self.view = v
        self.hview = None
        if self.hash_capacity:
            h = _lib.HashView()
            for s in range(self.num_shards):
                base = self.slab.ptrs[self.shard_owner[s]] + self.woff[s]
                h.keys[s] = base
                h.vals[s] = base + self.hash_capacity * 4
                h.capacity[s] = self.hash_capacity
                h.begin[s] = self.ranges[s][0]
            h.begin[self.num_shards] = self.ranges[-1][1]
            h.num_shards = self.num_shards
            self.hview = h
            self.overflow = torch.zeros(1, dtype=torch.int32, device=f"cuda:{self.device}")
            for s in self.own_shards:
                base = self.slab.local_ptr + self.woff[s]
                _lib.check(self.lib.mps_hash_clear(C.c_void_p(base), C.c_void_p(base + self.hash_capacity * 4),
                                                   self.hash_capacity, _lib.stream_ptr()), "hash_clear")
            torch.cuda.synchronize(self.device)
            control.barrier()
        self._timeout_s = float(os.environ.get("MINIPS_SPIN_TIMEOUT_S", "10"))
        self.num_workers = 1
        self._clock_view = None
        self.barrier_epoch = 0
        self.lock = threading.Condition()
        self.local_clock: Dict[int, int] = {}     # worker -> enqueued clock (host mirror)
        self.local_events: Dict[int, torch.cuda.Event] = {}
        self.local_workers: List[int] = []
        self.ckpt_barrier = threading.Barrier(1)
        self.write_epoch = 0      # bumped by every local write that bypasses a fused kernel's replica
        self.xchg = None          # symmetric scratch of the LR exchange kernel (models/lr.py), lazily
        self.bucket = None        # scratch of the bucketed sparse LR step, lazily

    # ---- views -------------------------------------------------------------------------
    def clock_view(self) -> _lib.ClockView:
        if self._clock_view is None:
            c = _lib.ClockView()
            for r in range(self.world):
                c.pad[r] = self.slab.ptrs[r] + self.pad_off
            c.num_ranks = self.world
            c.my_rank = self.rank
            c.num_workers = self.num_workers
            c.num_shards = self.num_shards
            c.own_first = self.own_first
            c.own_count = self.own_count
            c.model = {ModelType.SSP: _lib.MODEL_SSP, ModelType.BSP: _lib.MODEL_BSP,
                       ModelType.ASP: _lib.MODEL_ASP}[self.model_type]
            c.staleness = self.staleness
            c.timeout_cycles = int(self._timeout_s * 1.9e9)
            self._clock_view = c
        return self._clock_view

    def set_timeout(self, seconds: float):
        self._timeout_s = seconds
        self._clock_view = None

    def shard_tensor(self, s: int, staging: bool = False) -> torch.Tensor:
        """Local shard as a torch tensor (owner only)."""
        assert self.shard_owner[s] == self.rank
        b, e = self.ranges[s]
        off = self.soff[s] if staging else self.woff[s]
        return self.slab.local_tensor(off, (e - b) * self.k, self.dtype)

    def hash_items(self, s: int):
        """Live (keys, values) of a LOCAL hash shard, sorted by key."""
        assert self.hash_capacity and self.shard_owner[s] == self.rank
        dev = f"cuda:{self.device}"
        keys = torch.empty(self.hash_capacity, dtype=torch.int32, device=dev)
        vals = torch.empty(self.hash_capacity, dtype=torch.float32, device=dev)
        count = torch.zeros(1, dtype=torch.int32, device=dev)
        base = self.slab.local_ptr + self.woff[s]
        _lib.check(self.lib.mps_hash_export(C.c_void_p(base), C.c_void_p(base + self.hash_capacity * 4),
                                            self.hash_capacity, C.c_void_p(keys.data_ptr()),
                                            C.c_void_p(vals.data_ptr()),
                                            C.c_void_p(count.data_ptr()), _lib.stream_ptr()), "hash_export")
        n = int(count.item())
        k64 = keys[:n].to(torch.int64) & 0xFFFFFFFF
        order = torch.argsort(k64)
        return k64[order], vals[:n][order]

    def pad_tensor(self) -> torch.Tensor:
        return self.slab.local_tensor(self.pad_off, _lib.PAD_INTS, torch.int32)

    # ---- tracker lifecycle ---------------------------------------------------------------
    def reset_workers(self, num_workers: int, local_workers: Sequence[int], start_clock: int = 0):
        """ResetWorker (reference server/consistency/*_model.cpp ResetWorker, engine.cpp:208-244):
        re-initialise the progress tracker replica of THIS rank.  Callers barrier afterwards."""
        assert 0 < num_workers <= _lib.MAX_WORKERS
        self.num_workers = num_workers
        self._clock_view = None
        cv = self.clock_view()
        _lib.check(self.lib.mps_reset_pad(C.byref(cv), start_clock, _lib.stream_ptr()), "reset_pad")
        torch.cuda.synchronize(self.device)
        with self.lock:
            self.local_workers = list(local_workers)
            self.local_clock = {w: start_clock for w in local_workers}
            self.local_events = {}
            self.fold_clock = start_clock
            self.write_epoch += 1
            self.ckpt_barrier = threading.Barrier(max(1, len(self.local_workers)))

    def check_error(self):
        if self.hview is not None and int(self.overflow.item()) != 0:
            raise RuntimeError(f"table {self.table_id}: a hash shard overflowed (capacity {self.hash_capacity}); "
                               "create the table with a larger map_capacity")
        if self.lib.mps_kmeans_tc_error() > 0 or self.lib.mps_dense_lr_tc_error() > 0:
            raise RuntimeError("a tcgen05 kernel's mbarrier wait expired (pipeline protocol error); its results "
                               "are invalid")
        if int(self.pad_tensor()[_lib.PAD_ERROR].item()) != 0:
            raise RuntimeError(f"table {self.table_id}: device-side wait timed out "
                               f"(a peer did not reach the expected clock within {self._timeout_s}s)")

    def close(self):
        if self.xchg is not None:
            self.xchg.close()
            self.xchg = None
        if self.bucket is not None:
            self.bucket.close()
            self.bucket = None
        self.slab.close()

    # ---- whole-table helpers (host convenience, not hot path) ----------------------------
    def gather_full(self) -> torch.Tensor:
        """Dense copy of the whole table on this device (no consistency gate)."""
        out = torch.empty(self.num_keys * self.k, dtype=self.dtype, device=f"cuda:{self.device}")
        _lib.check(self.lib.mps_pull_dense(C.byref(self.view), None, 0, 0, C.c_void_p(out.data_ptr()),
                                           _lib.stream_ptr()), "pull_dense")
        return out


class KVClientTable:
    """Worker-side handle (reference worker/kv_client_table.hpp:45-71): Get / Add / Clock /
    CheckPoint for one (worker, table).  Every call only ENQUEUES device work on the
    worker's stream; consistency waits happen inside the kernels."""

    def __init__(self, table: DeviceTable, worker_id: int, thread_id: int,
                 stream: Optional[torch.cuda.Stream] = None, engine=None):
        self.table = table
        self.worker_id = worker_id
        self.thread_id = thread_id
        self.stream = stream
        self.engine = engine
        self.lib = table.lib
        self.progress = table.local_clock.get(worker_id, 0)
        self.dev = f"cuda:{table.device}"

    # -- helpers
    def _sp(self):
        return _lib.stream_ptr(self.stream)

    def _local_gate(self):
        """Order this worker's stream after the clocks of the OTHER local workers it depends
        on, so a same-GPU dependency is never an in-kernel spin (kernels of different streams
        are not guaranteed to co-run)."""
        t = self.table
        if len(t.local_workers) <= 1 or t.model_type == ModelType.ASP:
            return
        need = self.progress if t.model_type == ModelType.BSP else self.progress - t.staleness
        with t.lock:
            t.lock.wait_for(lambda: all(t.local_clock[w] >= need for w in t.local_workers
                                        if w != self.worker_id))
            evs = [t.local_events[w] for w in t.local_workers if w != self.worker_id and w in t.local_events]
        s = self.stream or torch.cuda.current_stream()
        for ev in evs:
            s.wait_event(ev)

    # -- API
    def get(self, keys, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        t = self.table
        k32 = keys_as_i32(keys, self.dev)
        n = k32.numel()
        if out is None:
            out = torch.empty(n * t.k, dtype=t.dtype, device=self.dev)
        self._local_gate()
        cv = t.clock_view()
        if t.hview is not None:
            _lib.check(self.lib.mps_hash_get(C.byref(t.hview), C.byref(cv), self.progress, C.c_void_p(k32.data_ptr()),
                                             C.c_void_p(out.data_ptr()), n, self._sp()), "hash_get")
            return out
        with torch.cuda.stream(self.stream):
            _lib.check(self.lib.mps_pull(C.byref(t.view), C.byref(cv), self.worker_id, self.progress,
                                         C.c_void_p(k32.data_ptr()), C.c_void_p(out.data_ptr()), n,
                                         self._sp()), "pull")
        return out

    def add(self, keys, vals: torch.Tensor):
        t = self.table
        k32 = keys_as_i32(keys, self.dev)
        v = torch.as_tensor(vals).to(self.dev, dtype=t.dtype, non_blocking=True).contiguous()
        assert v.numel() == k32.numel() * t.k
        t.write_epoch += 1
        if t.hview is not None:
            _lib.check(self.lib.mps_hash_add(C.byref(t.hview), C.c_void_p(k32.data_ptr()), C.c_void_p(v.data_ptr()),
                                             k32.numel(), C.c_void_p(t.overflow.data_ptr()), self._sp()), "hash_add")
            return
        _lib.check(self.lib.mps_push(C.byref(t.view), C.c_void_p(k32.data_ptr()), C.c_void_p(v.data_ptr()),
                                     k32.numel(), self._sp()), "push")
        if self.stream is not None:
            k32.record_stream(self.stream)
            v.record_stream(self.stream)

    def get_all(self, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Get of every key of the table (k-means centroids, LR eval) as slab copies."""
        t = self.table
        assert t.hview is None, "whole-table Get needs dense (Vector) storage"
        if out is None:
            out = torch.empty(t.num_keys * t.k, dtype=t.dtype, device=self.dev)
        self._local_gate()
        cv = t.clock_view()
        _lib.check(self.lib.mps_pull_dense(C.byref(t.view), C.byref(cv), self.worker_id, self.progress,
                                           C.c_void_p(out.data_ptr()), self._sp()), "pull_dense")
        return out

    def add_all(self, delta: torch.Tensor, zero_src: bool = False):
        t = self.table
        assert delta.is_cuda and delta.dtype == t.dtype and delta.numel() == t.num_keys * t.k
        t.write_epoch += 1
        _lib.check(self.lib.mps_push_dense(C.byref(t.view), C.c_void_p(delta.data_ptr()), int(zero_src),
                                           t.own_first, self._sp()), "push_dense")

    def clock(self):
        t = self.table
        new_clock = self.progress + 1
        fold = 0
        s = self.stream or torch.cuda.current_stream()
        # decide-who-folds, enqueue and publish the host mirror atomically: two local workers
        # clocking concurrently must not both conclude "the other one is last"
        with t.lock:
            others = [w for w in t.local_workers if w != self.worker_id]
            last_local = all(t.local_clock[w] >= new_clock for w in others)
            if t.model_type == ModelType.BSP and t.own_count > 0 and last_local:
                fold = 1
                for w in others:
                    if w in t.local_events:
                        s.wait_event(t.local_events[w])
            cv = t.clock_view()
            _lib.check(self.lib.mps_clock(C.byref(t.view), C.byref(cv), self.worker_id, new_clock, fold,
                                          self._sp()), "clock")
            ev = torch.cuda.Event()
            ev.record(s)
            t.local_clock[self.worker_id] = new_clock
            t.local_events[self.worker_id] = ev
            if fold:
                t.fold_clock = new_clock
            t.lock.notify_all()
        self.progress = new_clock

    def straggle(self, seconds: float):
        """Injected straggler: stall this worker's stream for `seconds` on the device."""
        _lib.check(self.lib.mps_straggle(int(seconds * 1e9), self._sp()), "straggle")

    def checkpoint(self):
        """Blocking dump of every shard + tracker state (kv_client_table.hpp:136-154)."""
        if self.engine is None:
            raise RuntimeError("checkpoint needs an Engine")
        self.engine.checkpoint_table(self.table.table_id, self)
