"""Keep a rank's host side next to its GPU: CPU affinity and the memory policy of the process are set
to the NUMA node the GPU's PCIe root hangs off, BEFORE the pinned batch buffers are allocated, so
every rank streams its batches from local DRAM instead of across the socket interconnect (with 8
ranks on a two-socket host that is the difference between each rank getting its own PCIe rate and
all of them sharing the inter-socket link).  Best effort: anything that is not there (sysfs entry,
permission, syscall) leaves the process as it was.  MINIPS_NUMA=0 switches it off."""
from __future__ import annotations

import ctypes
import os
from typing import Optional

_SYS_SET_MEMPOLICY = 238      # x86_64
_MPOL_PREFERRED = 1


def _read(path: str) -> Optional[str]:
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def _parse_cpulist(s: str):
    cpus = set()
    for part in s.split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_numa_node(device_index: int) -> Optional[int]:
    try:
        import torch
        props = torch.cuda.get_device_properties(device_index)
        bus = f"{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
    except Exception:  # noqa: BLE001
        return None
    node = _read(f"/sys/bus/pci/devices/{bus}/numa_node")
    if node is None:
        return None
    try:
        n = int(node)
    except ValueError:
        return None
    return n if n >= 0 else None


def bind_to_gpu_node(device_index: int) -> dict:
    """Returns what was done: {"node": n or None, "cpus": count or None, "mempolicy": bool}."""
    out = {"node": None, "cpus": None, "mempolicy": False}
    if os.environ.get("MINIPS_NUMA", "1") == "0":
        return out
    node = gpu_numa_node(device_index)
    if node is None:
        return out
    out["node"] = node
    cpulist = _read(f"/sys/devices/system/node/node{node}/cpulist")
    if cpulist:
        try:
            want = _parse_cpulist(cpulist) & set(os.sched_getaffinity(0))
            if want:
                os.sched_setaffinity(0, want)
                out["cpus"] = len(want)
        except (OSError, ValueError):
            pass
    try:
        libc = ctypes.CDLL(None, use_errno=True)
        mask = ctypes.c_ulong(1 << node)
        rc = libc.syscall(_SYS_SET_MEMPOLICY, _MPOL_PREFERRED, ctypes.byref(mask), ctypes.c_ulong(64))
        out["mempolicy"] = rc == 0
    except Exception:  # noqa: BLE001
        pass
    return out
