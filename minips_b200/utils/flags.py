"""Flag registry + Context with the reference's flag names.

The reference defines gflags per application and mirrors them into a Context singleton so that
library code can read them by name (base/context.cpp:13-61).  Same here: `define_*_flags`
register the reference's names and defaults, `parse` accepts the same `--name=value` syntax
(also `--name value`, `--noname`), and `Context` gives typed access with the reference's
"missing key reads as empty / 0 / false" behaviour.
"""
from __future__ import annotations

import sys
from typing import Dict, List, Optional


class Context:
    _inst: Optional["Context"] = None

    def __init__(self):
        self.values: Dict[str, str] = {}
        self.help: Dict[str, str] = {}
        self.iterations: Dict[int, int] = {}

    @classmethod
    def get_instance(cls) -> "Context":
        if cls._inst is None:
            cls._inst = Context()
        return cls._inst

    # -- registry
    def define(self, name: str, default, help: str = ""):
        if isinstance(default, bool):
            default = "true" if default else "false"
        self.values.setdefault(name, str(default))
        self.help[name] = help

    def parse(self, argv: Optional[List[str]] = None) -> List[str]:
        argv = list(sys.argv[1:] if argv is None else argv)
        positional = []
        i = 0
        while i < len(argv):
            a = argv[i]
            i += 1
            if not a.startswith("-") or len(a) < 2:
                positional.append(a)
                continue
            body = a.lstrip("-")
            if "=" in body:
                name, value = body.split("=", 1)
            else:
                name = body
                cur = self.values.get(name)
                if cur in ("true", "false"):
                    value = "true"
                elif name.startswith("no") and self.values.get(name[2:]) in ("true", "false"):
                    name, value = name[2:], "false"
                elif i < len(argv) and not argv[i].startswith("-"):
                    value = argv[i]
                    i += 1
                else:
                    value = "true"
            self.set(name, value)
        return positional

    # -- access (context.cpp:24-61)
    def set(self, name: str, value):
        if isinstance(value, bool):
            value = "true" if value else "false"
        value = str(value)
        if value in ("True", "TRUE"):
            value = "true"
        if value in ("False", "FALSE"):
            value = "false"
        self.values[name] = value

    def get_string(self, name: str) -> str:
        return self.values.get(name, "")

    def get_int32(self, name: str) -> int:
        try:
            return int(float(self.values.get(name, "0") or 0))
        except ValueError:
            return 0

    get_int64 = get_int32

    def get_double(self, name: str) -> float:
        try:
            return float(self.values.get(name, "0") or 0)
        except ValueError:
            return 0.0

    def get_bool(self, name: str) -> bool:
        return self.values.get(name, "") in ("true", "1")

    def get_iteration(self, worker_id: int) -> int:
        return self.iterations.get(worker_id, 0)

    def set_iteration(self, worker_id: int, it: int):
        self.iterations[worker_id] = it

    def usage(self) -> str:
        return "\n".join(f"  --{k}={v}    {self.help.get(k, '')}" for k, v in sorted(self.values.items()))

    def clear(self):
        self.values.clear()
        self.help.clear()
        self.iterations.clear()


def define_common_flags(ctx: Context):
    d = ctx.define
    d("my_id", 0, "node id of this process (defaults to RANK under torchrun)")
    d("config_file", "", "hostfile: one id:hostname:port per line")
    d("hdfs_namenode", "localhost", "kept for flag parity; hdfs:// paths map to $MINIPS_HDFS_ROOT")
    d("hdfs_namenode_port", 9000)
    d("assigner_master_port", 19201)
    d("input", "", "libsvm file or directory; empty = synthetic data generated in HBM")
    d("kModelType", "SSP", "ASP / SSP / BSP")
    d("kStorageType", "Vector", "Map / Vector")
    d("kStaleness", 0)
    d("num_dims", 0)
    d("batch_size", 1)
    d("num_iters", 1000)
    d("num_workers_per_node", 1)
    d("num_servers_per_node", 1)
    d("num_local_load_thread", 2)
    d("alpha", 0.1)
    d("with_injected_straggler", 0.0, "probability that an iteration sleeps U(0,100) ms")
    d("init_dump", False)
    d("use_weight_file", False, "resume from the last checkpoint")
    d("checkpoint_toggle", False)
    d("checkpoint_interval", 100)
    d("checkpoint_file_prefix", "")
    d("checkpoint_raw_prefix", "")
    d("weight_file_prefix", "", "unused (reference parity)")
    d("heartbeat_interval", -1, "<= 0: no master / heartbeats")
    d("relaunch_cmd", "")
    d("report_prefix", "")
    d("report_interval", -1)
    d("scale", False)
    d("scale_name", "")
    d("scale_port", 14563)
    d("scale_file", "")
    d("backend", "cuda", "cuda (device parameter server) or host (C++ actor runtime)")
    d("seed", 0)


def define_lr_flags(ctx: Context):
    define_common_flags(ctx)
    ctx.define("kSpeculation", 5, "extra precomputed batches (reference parity)")
    ctx.define("kSparseSSPRecorderType", "None", "unused (reference parity)")
    ctx.define("num_samples", 10000, "synthetic samples per node when --input is empty")
    ctx.define("nnz_per_row", 40)
    ctx.values["num_dims"] = ctx.values.get("num_dims") if ctx.values.get("num_dims", "0") != "0" else "20958"


def define_kmeans_flags(ctx: Context):
    define_common_flags(ctx)
    ctx.define("K", 2)
    ctx.define("kmeans_init_mode", "random", "random | kmeans++ | kmeans_parallel")
    ctx.define("report_worker", 0)
    ctx.define("num_nodes", 0, "unused (reference parity)")
    ctx.define("num_samples", 100000, "synthetic points per node when --input is empty")
    ctx.values["num_dims"] = ctx.values.get("num_dims") if ctx.values.get("num_dims", "0") != "0" else "123"
    ctx.values["batch_size"] = ctx.values.get("batch_size") if ctx.values.get("batch_size", "1") != "1" else "100"
    ctx.values["report_interval"] = "2" if ctx.values.get("report_interval") == "-1" else ctx.values["report_interval"]
