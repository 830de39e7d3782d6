"""Checkpoint / restore of device tables in the reference's dump formats.

  <prefix>server_params_<id>    one line, "<local_idx>:<val> " for every non-zero entry of the
                                shard, local_idx = key - range.begin   (server/vector_storage.hpp:54-73)
  <prefix>server_progress_<id>  "min_clock:<v> <tid>:<v> ..."           (server/util/progress_tracker.hpp:68-85)
  <prefix>worker_config_<id>    "<worker_id>:<iteration> ..."           (lib/svm_dumper.hpp:51-66)

  <prefix>server_layout         (this backend) JSON: node ids and the [begin, end) range behind every
                                server_params file, written by the first node.  Restore compares it with
                                the current partition: a cluster that grew (scale-out) or shrank since the
                                dump re-shards while it reads, and a node that was not there at dump time
                                takes the clock and the iteration of the others (the reference clones the
                                progress onto the scale node's thread, progress_tracker.hpp:112-117).

<id> is the node id; a node with several server shards appends ".<k>" for shards k >= 1.  Values are
written with round-trip precision (the reference's 6 significant digits lose weights) and Restore
parses what Dump wrote (the reference's Restore truncates to int).  Clocks / iterations are rounded
to the checkpoint interval like the reference (checkpoints are taken every `interval` iterations).

On the device backend CheckPoint() is COLLECTIVE: every worker of the task calls it at the same
iteration (there are no server threads that could dump on a remote worker's request).
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import numpy as np
import torch

from ..ops import _lib


def _paths(prefix: str, node_id: int, k: int, kind: str) -> str:
    return f"{prefix}{kind}_{node_id}" + (f".{k}" if k else "")


def _resolve(path: str) -> str:
    """hdfs://host:port/a/b -> $MINIPS_HDFS_ROOT/a/b (same mapping as csrc/core/file_io.cpp)."""
    if path.startswith("hdfs://"):
        rest = path[7:]
        tail = rest[rest.find("/"):] if "/" in rest else "/"
        path = os.environ.get("MINIPS_HDFS_ROOT", "/tmp/minips_hdfs") + tail
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    return path


def round_to_interval(v: int, interval: int = 100) -> int:
    """Nearest multiple of the checkpoint interval, halves up — the host runtime's RoundToInterval
    (csrc/core/server.cpp, lround) and the reference's ::round (progress_tracker.hpp:64-85)."""
    import math
    return int(math.floor(v / interval + 0.5)) * interval if interval > 0 else v


def write_sparse_line(path: str, values: np.ndarray):
    """The VectorStorage::Dump format."""
    idx = np.flatnonzero(values)
    digits = 17 if values.dtype == np.float64 else 9    # round-trip precision of the table's type
    with open(_resolve(path), "w") as f:
        step = 1 << 18
        for s in range(0, idx.size, step):
            chunk = idx[s:s + step]
            f.write(" ".join(f"{i}:{values[i]:.{digits}g}" for i in chunk))
            f.write(" ")


def read_sparse_line(path: str, size: int, dtype=np.float32) -> np.ndarray:
    out = np.zeros(size, dtype=dtype)
    p = _resolve(path)
    if not os.path.exists(p):
        return out
    with open(p) as f:
        for tok in f.read().split():
            i, v = tok.split(":")
            out[int(i)] = float(v)
    return out


def dump_table(engine, table, client=None, prefix: Optional[str] = None, interval: Optional[int] = None,
               iteration_map: Optional[Dict[int, int]] = None):
    """Collective, blocking dump of one table: shard values, progress tracker, iteration map.

    Every worker of the task calls it at the same iteration.  With several workers in one process
    they first meet on a thread barrier (so all their pushes are enqueued), ONE of them runs the
    process-level protocol (device sync, rank barrier, files, rank barrier) and the others wait for
    it: rank barriers pair by call order, so they must be issued once per process."""
    n_local = len(table.local_workers)
    if n_local > 1:
        if table.ckpt_barrier.wait() != 0:
            table.ckpt_barrier.wait()       # the leader is writing; released when it is done
            return
    try:
        _dump_table_once(engine, table, prefix, interval, iteration_map)
    finally:
        if n_local > 1:
            table.ckpt_barrier.wait()


def _dump_table_once(engine, table, prefix, interval, iteration_map):
    prefix = prefix if prefix is not None else str(engine.context.get("checkpoint_file_prefix", ""))
    if interval is None:
        interval = int(engine.context.get("checkpoint_interval", 100))
    torch.cuda.synchronize(table.device)
    engine.control.barrier()  # every push of every worker has landed
    engine.run_dump_callback()
    node_id = engine.node.id
    for k, s in enumerate(table.own_shards):
        if table.hview is not None:
            keys, vals = table.hash_items(s)
            lo = table.ranges[s][0]
            with open(_resolve(_paths(prefix, node_id, k, "server_params")), "w") as f:
                f.write(" ".join(f"{int(kk) - lo}:{float(v):.9g}" for kk, v in zip(keys.tolist(), vals.tolist())))
                f.write(" ")
            continue
        # BSP: staging is empty between clock edges when called at an iteration boundary
        vals = table.shard_tensor(s).detach().cpu().numpy()
        write_sparse_line(_paths(prefix, node_id, k, "server_params"), vals)
    pad = table.pad_tensor().cpu().numpy()
    spec = engine.worker_spec
    with open(_resolve(_paths(prefix, node_id, 0, "server_progress")), "w") as f:
        prog = pad[_lib.PAD_PROGRESS:_lib.PAD_PROGRESS + table.num_workers]
        f.write(f"min_clock:{round_to_interval(int(prog.min()) if prog.size else 0, interval)} ")
        for w in range(table.num_workers):
            tid = spec.worker_to_thread[w] if spec is not None and w in spec.worker_to_thread else w
            f.write(f"{tid}:{round_to_interval(int(prog[w]), interval)} ")
    it_map = iteration_map if iteration_map is not None else engine.iteration_map
    with open(_resolve(_paths(prefix, node_id, 0, "worker_config")), "w") as f:
        for w, it in sorted(it_map.items()):
            f.write(f"{w}:{round_to_interval(int(it), interval)} ")
    if engine.rank_index == 0:
        write_layout(prefix, table_layout(engine, table))
    engine.control.barrier()


def table_layout(engine, table) -> dict:
    """Which node's which file holds which keys (the partition at dump time)."""
    node_ids = [n.id for n in engine.nodes]
    per_node: Dict[int, int] = {}
    shards = []
    for s in range(table.num_shards):
        node = node_ids[table.shard_owner[s]]
        k = per_node.get(node, 0)
        per_node[node] = k + 1
        b, e = table.ranges[s]
        shards.append({"node": int(node), "k": int(k), "begin": int(b), "end": int(e)})
    return {"nodes": [int(i) for i in node_ids], "vals_per_key": int(table.k), "shards": shards}


def write_layout(prefix: str, layout: dict):
    with open(_resolve(prefix + "server_layout"), "w") as f:
        json.dump(layout, f)


def read_layout(prefix: str) -> Optional[dict]:
    p = _resolve(prefix + "server_layout")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        return json.load(f)


def read_resharded(prefix: str, layout: dict, begin: int, end: int, k: int, dtype=np.float32,
                   cache: Optional[dict] = None) -> np.ndarray:
    """Values of keys [begin, end) collected from the dump files of a DIFFERENT partition."""
    out = np.zeros((end - begin) * k, dtype=dtype)
    cache = cache if cache is not None else {}
    for sh in layout["shards"]:
        b, e = max(begin, sh["begin"]), min(end, sh["end"])
        if b >= e:
            continue
        key = (sh["node"], sh["k"])
        if key not in cache:
            cache[key] = read_sparse_line(_paths(prefix, sh["node"], sh["k"], "server_params"),
                                          (sh["end"] - sh["begin"]) * k, dtype)
        src = cache[key]
        out[(b - begin) * k:(e - begin) * k] = src[(b - sh["begin"]) * k:(e - sh["begin"]) * k]
    return out


def _read_progress(prefix: str, node_id: int) -> Optional[int]:
    p = _resolve(_paths(prefix, node_id, 0, "server_progress"))
    if not os.path.exists(p):
        return None
    for tok in open(p).read().split():
        kname, v = tok.split(":")
        if kname == "min_clock":
            return int(v)
    return 0


def _read_config(prefix: str, node_id: int) -> Optional[Dict[int, int]]:
    p = _resolve(_paths(prefix, node_id, 0, "worker_config"))
    if not os.path.exists(p):
        return None
    out: Dict[int, int] = {}
    for tok in open(p).read().split():
        w, it = tok.split(":")
        out[int(w)] = int(it)
    return out


def restore_table(engine, table, prefix: Optional[str] = None) -> Dict[int, int]:
    """Collective restore of a table's shards + tracker; returns the iteration map.  The partition may
    differ from the one that was dumped (see server_layout above)."""
    prefix = prefix if prefix is not None else str(engine.context.get("checkpoint_file_prefix", ""))
    node_id = engine.node.id
    dev = f"cuda:{table.device}"
    assert table.hview is None, "restore of hash-backed tables goes through add()"
    np_dtype = np.float32 if table.dtype == torch.float32 else np.float64
    layout = read_layout(prefix)
    dumped = {(sh["node"], sh["k"], sh["begin"], sh["end"]) for sh in layout["shards"]} if layout else None
    same_cluster = layout is None or [int(n.id) for n in engine.nodes] == layout["nodes"]
    cache: dict = {}
    for k, s in enumerate(table.own_shards):
        b, e = table.ranges[s]
        if dumped is None or (int(node_id), k, int(b), int(e)) in dumped:
            vals = read_sparse_line(_paths(prefix, node_id, k, "server_params"), (e - b) * table.k, np_dtype)
        else:
            vals = read_resharded(prefix, layout, int(b), int(e), table.k, np_dtype, cache)
        table.shard_tensor(s).copy_(torch.from_numpy(vals).to(dev))
        table.write_epoch += 1
        if table.model_type.name == "BSP":
            table.shard_tensor(s, staging=True).zero_()
    clock = _read_progress(prefix, node_id)
    it_map = _read_config(prefix, node_id)
    if not same_cluster:
        # membership changed since the dump: every node resumes from the cluster-wide minimum, and workers
        # that did not exist then start there too (Engine.get_iteration falls back to default_iteration)
        old = layout["nodes"]
        clocks = [c for c in (_read_progress(prefix, n) for n in old) if c is not None]
        clock = min(clocks) if clocks else 0
        merged: Dict[int, int] = {}
        for n in old:
            merged.update(_read_config(prefix, n) or {})
        low = min(merged.values()) if merged else 0
        it_map = {w: low for w in merged}
        engine.default_iteration = low
    torch.cuda.synchronize(table.device)
    table.restored_clock = clock or 0
    engine.control.barrier()
    return it_map or {}
