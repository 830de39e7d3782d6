"""minips_b200 — a Blackwell-native parameter server with the capabilities of MiniPs.

Layout
  parallel/   Engine, MLTask, WorkerSpec, device tables, symmetric heap (the PS runtime)
  ops/        ctypes bindings of the hand-written sm_100a kernels (libminips_kernels.so)
  models/     sparse / dense logistic regression, mini-batch k-means
  utils/      flags (Context), hostfile parsing, checkpoint formats, timing, clock sampling
  apps/       CLI entry points mirroring the reference executables (BasicExample, LRExample, KMeans)
  core        pybind11 bindings of the C++ host runtime (csrc/core), loaded lazily
"""
from .parallel.types import (MLTask, ModelType, Node, SimpleIdMapper, StorageType, WorkerAlloc, WorkerSpec,
                             even_ranges)

__version__ = "0.1.0"

__all__ = ["Engine", "Info", "KVClientTable", "MLTask", "ModelType", "Node", "SimpleIdMapper", "StorageType",
           "WorkerAlloc", "WorkerSpec", "even_ranges"]


def __getattr__(name):
    # device-backend symbols import torch lazily so `import minips_b200` stays light
    if name in ("Engine", "Info"):
        from .parallel import engine as _e
        return getattr(_e, name)
    if name == "KVClientTable":
        from .parallel.device_table import KVClientTable
        return KVClientTable
    raise AttributeError(name)
