"""Symmetric heap: one slab per rank, mapped into every rank of the NVLink domain.

This is the data plane that replaces the reference's ZMQ mailbox for kAdd/kGet traffic
(reference comm/mailbox.cpp:106-129, 231-308): instead of sending a message to the node
that owns a key range, a kernel dereferences the owner's memory.  Handles (64-byte CUDA IPC
handles) are exchanged once over the control-plane process group.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from ..ops import _lib


class _ControlPlane:
    """all_gather_object over a (gloo) group, or the trivial single-process version."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.rank = dist.get_rank(group)
            self.world = dist.get_world_size(group)
        else:
            self.rank, self.world = 0, 1

    def all_gather(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj, group=self.group)
        return out

    def barrier(self):
        if self.world > 1:
            self.dist.barrier(group=self.group)


class SymmetricBuffer:
    """`nbytes` of zeroed device memory on every rank; ptrs[r] addresses rank r's copy from
    this rank's device (ptrs[rank] is the local allocation).

    multicast=True asks for a buffer that the NVSwitch also maps as ONE multicast address
    (`mc_ptr`, 0 when the fabric or driver has no NVLS): a `multimem.ld_reduce` there returns the
    sum over all ranks' copies, a `multimem.st` writes all of them.  That needs memory from the
    driver's virtual-memory API bound to a multicast object, which torch's symmetric-memory
    allocator provides (cuMemCreate + cuMulticastBindMem, handles exchanged over the process
    group's store); plain buffers stay cudaMalloc + CUDA IPC.
    """

    def __init__(self, nbytes: int, control: Optional[_ControlPlane] = None, device: Optional[int] = None,
                 multicast: bool = False):
        self.lib = _lib.load()
        self.control = control or _ControlPlane()
        self.rank, self.world = self.control.rank, self.control.world
        self.nbytes = max(256, (int(nbytes) + 255) // 256 * 256)
        self.device = torch.cuda.current_device() if device is None else device
        _lib.check(self.lib.mps_set_device(self.device), "set_device")
        self.mc_ptr = 0
        self._symm = None
        self._opened: List[int] = []
        self._closed = False
        if multicast and self.world > 1 and self._try_symm():
            return
        p = C.c_void_p()
        _lib.check(self.lib.mps_heap_alloc(self.nbytes, C.byref(p)), f"heap_alloc({self.nbytes})")
        self.local_ptr = p.value
        self.ptrs: List[int] = [0] * self.world
        self.ptrs[self.rank] = self.local_ptr
        if self.world > 1:
            h = C.create_string_buffer(64)
            _lib.check(self.lib.mps_ipc_get_handle(C.c_void_p(self.local_ptr), h), "ipc_get_handle")
            handles = self.control.all_gather(bytes(h.raw))
            for r, hb in enumerate(handles):
                if r == self.rank:
                    continue
                q = C.c_void_p()
                _lib.check(self.lib.mps_ipc_open_handle(C.c_char_p(hb), C.byref(q)), f"ipc_open_handle(rank {r})")
                self.ptrs[r] = q.value
                self._opened.append(q.value)
            self.control.barrier()

    def _try_symm(self) -> bool:
        """Allocate through torch's symmetric memory.  Every rank reports whether its local part
        worked before the collective rendezvous, so a rank that cannot allocate sends everybody to
        the IPC path instead of leaving the others inside a collective."""
        import os
        # collective decision: one rank that wants plain IPC sends everybody there
        if not all(self.control.all_gather(os.environ.get("MINIPS_HEAP", "symm") != "ipc")):
            return False
        dist = self.control.dist
        t, err = None, None
        try:
            import torch.distributed._symmetric_memory as symm
            t = symm.empty(self.nbytes, dtype=torch.uint8, device=f"cuda:{self.device}")
        except Exception as e:  # noqa: BLE001 - any failure means "use IPC"
            err = repr(e)
        if not all(e is None for e in self.control.all_gather(err)):
            return False
        try:
            hdl = symm.rendezvous(t, self.control.group if self.control.group is not None else dist.group.WORLD)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            mc = int(hdl.multicast_ptr) if hdl.has_multicast_support else 0
        except Exception as e:  # noqa: BLE001
            err = repr(e)
        if not all(e is None for e in self.control.all_gather(err)):
            return False
        t.zero_()
        torch.cuda.synchronize(self.device)
        self._symm = (t, hdl)
        self.ptrs = ptrs
        self.local_ptr = ptrs[self.rank]
        assert self.local_ptr == t.data_ptr()
        # all ranks must agree on whether the multicast mapping exists
        self.mc_ptr = mc if all(self.control.all_gather(mc != 0)) else 0
        self.control.barrier()
        return True

    def local_tensor(self, offset_bytes: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """Zero-copy torch view of a region of the LOCAL slab."""
        itemsize = torch.empty((), dtype=dtype).element_size()
        assert offset_bytes + numel * itemsize <= self.nbytes
        return _from_ptr(self.local_ptr + offset_bytes, numel, dtype, self.device, self)

    def close(self):
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
        except Exception:
            pass
        if self._symm is not None:
            self._symm = None  # the allocator unmaps when the tensor and its handle go away
            return
        for q in self._opened:
            self.lib.mps_ipc_close_handle(C.c_void_p(q))
        if self.world > 1:
            try:
                self.control.barrier()
            except Exception:
                pass
        self.lib.mps_heap_free(C.c_void_p(self.local_ptr))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch.as_tensor can alias raw memory."""

    def __init__(self, ptr, numel, typestr, owner):
        self.__cuda_array_interface__ = {
            "shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 3, "strides": None,
        }
        self._owner = owner


_TYPESTR = {torch.float32: "<f4", torch.float64: "<f8", torch.int32: "<i4", torch.int64: "<i8",
            torch.uint8: "|u1", torch.bfloat16: None}


def _from_ptr(ptr: int, numel: int, dtype: torch.dtype, device: int, owner) -> torch.Tensor:
    if numel == 0:
        return torch.empty(0, dtype=dtype, device=f"cuda:{device}")
    ts = _TYPESTR[dtype]
    t = torch.as_tensor(_CudaArray(ptr, numel, ts, owner), device=f"cuda:{device}")
    t._minips_owner = owner  # keep the slab alive as long as the view is
    return t
