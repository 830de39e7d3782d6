"""Tracing helpers: NVTX ranges and a per-phase CUDA-event timer.

The reference times itself with ad-hoc std::chrono stamps in the apps (lr_example.cpp:238,357,
376-377; kmeans.cpp:221,278-291) and has no tracer.  Here:

* `nvtx_range("name")` brackets host code with an NVTX range (visible in nsys / ncu --nvtx) when
  `MINIPS_NVTX=1`; otherwise it is a no-op that costs one branch, so it can stay in hot loops.
* `PhaseTimer` records CUDA events around named phases on the current stream and reports, per
  phase, the total device time as the MAX over ranks (the number a lock-step job is gated by).
* the dense LR kernel additionally carries its own in-kernel phase timeline
  (`mps_lr_train_dense_ex(..., timeline)`, see `bench.py --timeline`).
"""
from __future__ import annotations

import contextlib
import os
from collections import OrderedDict
from typing import Dict, List, Tuple

import torch

_NVTX = os.environ.get("MINIPS_NVTX", "0") == "1"


@contextlib.contextmanager
def nvtx_range(name: str):
    if _NVTX and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


def traced(name: str):
    """Decorator form of nvtx_range; returns the function untouched when NVTX is off."""
    def deco(fn):
        if not _NVTX:
            return fn
        import functools

        @functools.wraps(fn)
        def wrapper(*a, **k):
            with nvtx_range(name):
                return fn(*a, **k)
        return wrapper
    return deco


def nvtx_mark(name: str):
    if _NVTX and torch.cuda.is_available():
        torch.cuda.nvtx.mark(name)


class PhaseTimer:
    """CUDA-event timer for named phases of an iteration.

        pt = PhaseTimer()
        for ...:
            with pt.phase("pull"): ...
            with pt.phase("compute"): ...
        pt.summary(device)   # {"pull": {"ms": total, "calls": n, "ms_max_over_ranks": ...}, ...}

    Events are only resolved in summary(), so recording does not synchronise the stream.
    """

    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self._spans: "OrderedDict[str, List[Tuple[torch.cuda.Event, torch.cuda.Event]]]" = OrderedDict()

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with nvtx_range(name):
            e0.record()
            try:
                yield
            finally:
                e1.record()
                self._spans.setdefault(name, []).append((e0, e1))

    def summary(self, device=None, reduce_over_ranks: bool = True) -> Dict[str, Dict[str, float]]:
        out: Dict[str, Dict[str, float]] = OrderedDict()
        if not self.enabled:
            return out
        torch.cuda.synchronize()
        for name, spans in self._spans.items():
            total = sum(a.elapsed_time(b) for a, b in spans)
            rec = {"ms": total, "calls": len(spans)}
            if reduce_over_ranks:
                from .benchkit import max_over_ranks
                rec["ms_max_over_ranks"] = max_over_ranks(total, device if device is not None else "cuda")
            out[name] = rec
        return out

    def reset(self):
        self._spans.clear()
