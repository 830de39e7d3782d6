"""Driver-level value types shared by the host and device backends.

Mirrors the reference's public driver vocabulary:
  Node            base/node.hpp:8-22
  ModelType/StorageType   driver/engine.hpp:25-30
  WorkerAlloc / MLTask    driver/ml_task.hpp:12-66
  SimpleIdMapper  driver/simple_id_mapper.hpp:85-95, .cpp:13-89 (thread-id layout)
  WorkerSpec      driver/worker_spec.cpp:11-58
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple


class ModelType(enum.IntEnum):
    SSP = 0
    BSP = 1
    ASP = 2


class StorageType(enum.IntEnum):
    Map = 0
    Vector = 1


@dataclass(frozen=True)
class Node:
    id: int
    hostname: str = "localhost"
    port: int = 0
    is_master: bool = False

    def debug_string(self) -> str:
        return f"{{ id: {self.id} hostname: {self.hostname} port: {self.port} }}"


@dataclass
class WorkerAlloc:
    node_id: int
    num_workers: int


Range = Tuple[int, int]  # half-open [begin, end), base/third_party/range.h:14-29


class SimpleIdMapper:
    """tid / 1000 = node; servers node*1000+[0,50); worker helper node*1000+50; user workers
    node*1000+[100,1000)."""

    kMaxNodeId = 1000
    kMaxThreadsPerNode = 1000
    kMaxBgThreadsPerNode = 100
    kWorkerHelperThreadId = 50

    def __init__(self, node: Node, nodes: Sequence[Node]):
        self.node = node
        self.nodes = list(nodes)
        self.node2server: Dict[int, List[int]] = {}
        self.node2worker_helper: Dict[int, List[int]] = {}
        self.node2worker: Dict[int, set] = {}

    def init(self, num_server_threads_per_node: int = 1):
        assert 0 < num_server_threads_per_node <= self.kWorkerHelperThreadId
        self.node2server.clear()
        self.node2worker_helper.clear()
        self.node2worker.clear()
        for n in self.nodes:
            assert n.id < self.kMaxNodeId
            self.node2server[n.id] = [n.id * self.kMaxThreadsPerNode + i for i in range(num_server_threads_per_node)]
            self.node2worker_helper[n.id] = [n.id * self.kMaxThreadsPerNode + self.kWorkerHelperThreadId]
            self.node2worker[n.id] = set()

    def get_node_id_for_thread(self, tid: int) -> int:
        return tid // self.kMaxThreadsPerNode

    def allocate_worker_thread(self, node_id: int) -> int:
        assert node_id in self.node2worker_helper
        used = self.node2worker[node_id]
        assert len(used) < self.kMaxThreadsPerNode - self.kMaxBgThreadsPerNode
        for i in range(self.kMaxBgThreadsPerNode, self.kMaxThreadsPerNode):
            tid = i + node_id * self.kMaxThreadsPerNode
            if tid not in used:
                used.add(tid)
                return tid
        raise RuntimeError("no free worker thread id")

    def deallocate_worker_thread(self, node_id: int, tid: int):
        self.node2worker[node_id].remove(tid)

    def get_server_threads_for_id(self, node_id: int) -> List[int]:
        return list(self.node2server[node_id])

    def get_worker_helper_threads_for_id(self, node_id: int) -> List[int]:
        return list(self.node2worker_helper[node_id])

    def get_worker_threads_for_id(self, node_id: int) -> List[int]:
        return sorted(self.node2worker[node_id])

    def get_all_server_threads(self) -> List[int]:
        out: List[int] = []
        for n in self.nodes:
            out.extend(self.node2server[n.id])
        return out


class WorkerSpec:
    """Task-local worker ids 0..W-1 in worker_alloc order, worker<->node<->thread maps."""

    def __init__(self, worker_alloc: Sequence[WorkerAlloc]):
        self.worker_to_node: Dict[int, int] = {}
        self.worker_to_thread: Dict[int, int] = {}
        self.thread_to_worker: Dict[int, int] = {}
        self.node_to_workers: Dict[int, List[int]] = {}
        self.node_to_threads: Dict[int, List[int]] = {}
        self.thread_ids: set = set()
        self.num_workers = 0
        for a in worker_alloc:
            assert a.node_id not in self.node_to_workers, "duplicate node in worker_alloc"
            for _ in range(a.num_workers):
                self.worker_to_node[self.num_workers] = a.node_id
                self.node_to_workers.setdefault(a.node_id, []).append(self.num_workers)
                self.num_workers += 1

    def has_local_workers(self, node_id: int) -> bool:
        return node_id in self.node_to_workers

    def get_local_workers(self, node_id: int) -> List[int]:
        return list(self.node_to_workers[node_id])

    def get_local_threads(self, node_id: int) -> List[int]:
        return list(self.node_to_threads[node_id])

    def get_all_thread_ids(self) -> List[int]:
        return sorted(self.thread_ids)

    def insert_worker_id_thread_id(self, worker_id: int, thread_id: int):
        assert worker_id in self.worker_to_node
        assert worker_id not in self.worker_to_thread
        assert thread_id not in self.thread_to_worker
        self.worker_to_thread[worker_id] = thread_id
        self.thread_to_worker[thread_id] = worker_id
        self.node_to_threads.setdefault(self.worker_to_node[worker_id], []).append(thread_id)
        self.thread_ids.add(thread_id)


@dataclass
class MLTask:
    func: Optional[Callable] = None
    worker_alloc: List[WorkerAlloc] = field(default_factory=list)
    tables: List[int] = field(default_factory=list)

    def set_lambda(self, func: Callable):
        self.func = func

    def run_lambda(self, info):
        return self.func(info)

    def set_worker_alloc(self, worker_alloc):
        self.worker_alloc = [a if isinstance(a, WorkerAlloc) else WorkerAlloc(*a) for a in worker_alloc]

    def get_worker_alloc(self):
        return self.worker_alloc

    def set_tables(self, tables):
        self.tables = list(tables)

    def get_tables(self):
        return self.tables

    def is_setup(self) -> bool:
        return self.func is not None and bool(self.worker_alloc) and bool(self.tables)


def even_ranges(num_dims: int, num_shards: int) -> List[Range]:
    """Engine::getRanges (driver/engine.hpp:158-172): num_dims / S each, last takes the rest."""
    w = num_dims // num_shards
    out = [(w * i, w * (i + 1)) for i in range(num_shards - 1)]
    out.append((w * (num_shards - 1), num_dims))
    return out
