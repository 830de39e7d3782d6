"""Engine — per-process facade of the device backend (reference driver/engine.hpp:42-378).

One process per GPU ("node").  StartEverything brings up the control plane and the id
mapper; CreateTable allocates the table's shards in the symmetric heap; Run executes the
task lambda on this node's workers between two barriers, exactly like the reference
(driver/engine.cpp:246-297) — except that there are no server / sender / receiver threads:
Get/Add/Clock are kernels that address the owning GPU's shard directly.
"""
from __future__ import annotations

import os
import threading
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence

import torch

from .device_table import DeviceTable, KVClientTable
from .heap import _ControlPlane
from ..utils.tracing import traced
from .types import (MLTask, ModelType, Node, Range, SimpleIdMapper, StorageType, WorkerAlloc, WorkerSpec,
                    even_ranges)


@dataclass
class Info:
    """Per-worker context handed to the task lambda (reference driver/info.hpp:17-43)."""
    thread_id: int
    worker_id: int
    engine: "Engine"
    stream: Optional[torch.cuda.Stream] = None
    local_id: int = 0

    def create_kv_client_table(self, table_id: int) -> KVClientTable:
        t = self.engine.tables[table_id]
        return KVClientTable(t, self.worker_id, self.thread_id, self.stream, self.engine)

    # reference spelling
    CreateKVClientTable = create_kv_client_table

    def debug_string(self) -> str:
        return f"thread_id: {self.thread_id} worker_id: {self.worker_id}"


class Engine:
    def __init__(self, node: Optional[Node] = None, nodes: Optional[Sequence[Node]] = None,
                 master_node: Optional[Node] = None, device: Optional[int] = None, group=None):
        import torch.distributed as dist
        self.dist = dist
        if nodes is None:
            world = dist.get_world_size() if dist.is_initialized() else 1
            rank = dist.get_rank() if dist.is_initialized() else 0
            nodes = [Node(i, "localhost", 0) for i in range(world)]
            node = nodes[rank]
        self.node = node
        self.nodes = sorted(nodes, key=lambda n: n.id)
        self.master_node = master_node
        self.device = device if device is not None else (
            torch.cuda.current_device() if torch.cuda.is_available() else 0)
        self._group = group
        self.control: Optional[_ControlPlane] = None
        self.id_mapper: Optional[SimpleIdMapper] = None
        self.tables: Dict[int, DeviceTable] = {}
        self.model_count = 0
        self.num_server_threads_per_node = 1
        self.worker_spec: Optional[WorkerSpec] = None
        self.dump_callback: Optional[Callable] = None
        self.context: Dict[str, object] = {}
        self.iteration_map: Dict[int, int] = {}
        self.default_iteration = 0      # where a worker that is in no checkpoint starts (set by a re-sharding restore)
        self.started = False
        self.numa = None
        self.rank_index = 0

    # ---- lifecycle (engine.cpp:17-44) -------------------------------------------------------
    def start_everything(self, num_server_threads_per_node: int = 1):
        dist = self.dist
        if dist.is_initialized() and dist.get_world_size() > 1 and self._group is None:
            # control plane rides on gloo so it never contends with kernels for the GPU
            self._group = dist.new_group(backend="gloo")
        self.control = _ControlPlane(self._group)
        assert self.control.world == len(self.nodes), "one process per node expected"
        self.rank = self.control.rank
        assert self.nodes[self.rank].id == self.node.id, "node ids must be in rank order"
        self.rank_index = self.rank
        self.num_server_threads_per_node = num_server_threads_per_node
        self.id_mapper = SimpleIdMapper(self.node, self.nodes)
        self.id_mapper.init(num_server_threads_per_node)
        if torch.cuda.is_available():
            torch.cuda.set_device(self.device)
        self.started = True

    def stop_everything(self):
        for t in self.tables.values():
            t.close()
        self.tables.clear()
        self.started = False

    StartEverything = start_everything
    StopEverything = stop_everything

    def force_quit(self, leaving: bool) -> bool:
        """Elastic shrink: nodes with `leaving` set leave the cluster before training (the reference's
        data-less node broadcasts kForceQuit, lr_example.cpp:145-152; its peers drop it from the node list,
        the barrier accounting and the id mapper: comm/mailbox.cpp:159-171, driver/engine.cpp:33-36).

        Collective over the current nodes, called after start_everything and BEFORE any table exists.  The
        survivors get a control group, node list and id mapper over the remaining nodes, so every table
        created afterwards is partitioned over the survivors' HBM only.  Returns False on a leaving node:
        it calls stop_everything and returns, like the reference."""
        assert self.started and not self.tables, "membership changes before the tables are created"
        flags = [bool(x) for x in self.control.all_gather(bool(leaving))]
        if not any(flags):
            return True
        keep = [i for i, f in enumerate(flags) if not f]
        assert keep, "every node asked to leave"
        self.left_nodes = [self.nodes[i] for i, f in enumerate(flags) if f]
        if flags[self.rank]:
            self.started = False
            return False
        dist = self.dist
        sub = self._group
        if self.control.world > 1:
            ranks = [dist.get_global_rank(self._group, i) if self._group is not None else i for i in keep]
            # only the survivors create the group (the others are on their way out)
            sub = dist.new_group(ranks=ranks, backend="gloo", use_local_synchronization=True)
        self.nodes = [self.nodes[i] for i in keep]
        self._group = sub
        self.control = _ControlPlane(sub)
        self.rank = self.control.rank
        self.rank_index = self.rank
        self.id_mapper = SimpleIdMapper(self.node, self.nodes)
        self.id_mapper.init(self.num_server_threads_per_node)
        return True

    ForceQuit = force_quit

    # ---- tables (engine.hpp:289-357) ------------------------------------------------------------
    def get_ranges(self, num_dims: int) -> List[Range]:
        return even_ranges(num_dims, len(self.nodes) * self.num_server_threads_per_node)

    def create_table(self, ranges: Optional[Sequence[Range]] = None, model_type: ModelType = ModelType.SSP,
                     storage_type: StorageType = StorageType.Vector, staleness: int = 0, *,
                     num_dims: Optional[int] = None, dtype: torch.dtype = torch.float32,
                     vals_per_key: int = 1, map_capacity: Optional[int] = None) -> int:
        """Collective: call with identical arguments, in the same order, on every node."""
        assert self.started
        if ranges is None:
            assert num_dims is not None
            ranges = self.get_ranges(num_dims)
        server_tids = self.id_mapper.get_all_server_threads()
        assert len(ranges) == len(server_tids), "one range per server thread"
        node_ids = [n.id for n in self.nodes]
        owner = [node_ids.index(self.id_mapper.get_node_id_for_thread(t)) for t in server_tids]
        tid = self.model_count
        self.model_count += 1
        # StorageType.Map over a bounded range behaves exactly like Vector (missing keys read 0, adds
        # accumulate) and is stored densely unless map_capacity asks for real hash shards
        self.tables[tid] = DeviceTable(tid, ranges, owner, self.control, model_type, staleness, storage_type,
                                       dtype, vals_per_key, self.device,
                                       hash_capacity=map_capacity if storage_type == StorageType.Map else None)
        return tid

    CreateTable = create_table

    @traced("engine.barrier")
    def barrier(self):
        """Process barrier (reference Mailbox::Barrier via Engine::Barrier, engine.cpp:186-188)."""
        if torch.cuda.is_available():
            torch.cuda.synchronize(self.device)
        self.control.barrier()

    Barrier = barrier

    @traced("engine.device_barrier")
    def device_barrier(self, table_id: int, stream=None):
        """Barrier among the ranks ON THE DEVICE (flags in the table's signal pad, no host round
        trip): every rank's `stream` continues only after every rank enqueued its barrier.
        Collective — all ranks must call it the same number of times."""
        import ctypes as C
        from ..ops import _lib
        t = self.tables[table_id]
        t.barrier_epoch += 1
        cv = t.clock_view()
        _lib.check(t.lib.mps_barrier(C.byref(cv), t.barrier_epoch, _lib.stream_ptr(stream)), "barrier")

    # ---- run (engine.cpp:246-297) ---------------------------------------------------------------
    def allocate_workers(self, worker_alloc: Sequence[WorkerAlloc]) -> WorkerSpec:
        spec = WorkerSpec(worker_alloc)
        for a in worker_alloc:
            for w in spec.get_local_workers(a.node_id):
                spec.insert_worker_id_thread_id(w, self.id_mapper.allocate_worker_thread(a.node_id))
        return spec

    @traced("engine.run")
    def run(self, task: MLTask):
        assert task.is_setup(), "task needs lambda, worker_alloc and tables"
        spec = self.allocate_workers(task.get_worker_alloc())
        self.worker_spec = spec
        local = spec.get_local_workers(self.node.id) if spec.has_local_workers(self.node.id) else []
        for tid in task.get_tables():
            # a restored table keeps its clocks (ssp_model.cpp:98-103: ResetWorker leaves the tracker alone
            # under use_weight_file; progress_tracker.hpp:87-122 restores them)
            t = self.tables[tid]
            t.reset_workers(spec.num_workers, local, start_clock=int(getattr(t, "restored_clock", 0) or 0))
            t.restored_clock = 0
        self.barrier()
        errors: List[BaseException] = []
        if len(local) == 1:
            w = local[0]
            info = Info(spec.worker_to_thread[w], w, self, None, 0)
            try:
                task.run_lambda(info)
            except BaseException as e:  # noqa: BLE001 - re-raised after the barrier
                errors.append(e)
        elif len(local) > 1:
            threads = []
            for i, w in enumerate(local):
                info = Info(spec.worker_to_thread[w], w, self, torch.cuda.Stream(self.device), i)

                def body(info=info):
                    try:
                        torch.cuda.set_device(self.device)
                        with torch.cuda.stream(info.stream):
                            task.run_lambda(info)
                        info.stream.synchronize()
                    except BaseException as e:  # noqa: BLE001
                        errors.append(e)
                        for tid in task.get_tables():
                            with self.tables[tid].lock:
                                self.tables[tid].local_clock = {k: 1 << 30 for k in self.tables[tid].local_clock}
                                self.tables[tid].lock.notify_all()

                th = threading.Thread(target=body, name=f"worker-{w}")
                th.start()
                threads.append(th)
            for th in threads:
                th.join()
        else:
            self._serve_until_done(task)
        if local:
            torch.cuda.synchronize(self.device)
            self._signal_done()
        for w in local:
            self.id_mapper.deallocate_worker_thread(self.node.id, spec.worker_to_thread[w])
        for tid in task.get_tables():
            try:
                self.tables[tid].check_error()
            except RuntimeError as e:
                errors.append(e)
        self.barrier()  # "let on-the-fly messages be received" (engine.cpp:295-296)
        if errors:
            raise errors[0]

    Run = run

    # Ranks that own BSP shards but run no worker of the task still have to apply the clock
    # edges of their shards (the reference's server thread does it, bsp_model.cpp:14-32).
    def _signal_done(self):
        if self.control.world > 1:
            self._done_work = self.dist.barrier(group=self.control.group, async_op=True)
            self._done_work.wait()

    def _serve_until_done(self, task: MLTask):
        """This rank owns shards of the task's tables but runs none of its workers.  For every BSP table a
        resident device kernel (mps_fold_server) applies the clock edges as the workers' clocks arrive in
        this rank's pad; the host only waits for the end of the task and then tells the kernels to stop."""
        import ctypes as C
        from ..ops import _lib
        if self.control.world == 1:
            return
        done = self.dist.barrier(group=self.control.group, async_op=True)
        tabs = [self.tables[t] for t in task.get_tables()
                if self.tables[t].model_type == ModelType.BSP and self.tables[t].own_count > 0]
        flag = torch.zeros(1, dtype=torch.int32).pin_memory()
        streams = []
        for t in tabs:
            s = torch.cuda.Stream(self.device)        # one resident kernel per table, side by side
            start = int(t.pad_tensor()[_lib.PAD_FOLDED + t.own_first].item())
            _lib.check(t.lib.mps_fold_server(C.byref(t.view), C.byref(t.clock_view()), start,
                                             C.c_void_p(flag.data_ptr()), _lib.stream_ptr(s)), "fold_server")
            streams.append(s)
        done.wait()
        flag[0] = 1                                   # the kernels poll it through the pinned mapping
        for s in streams:
            s.synchronize()

    # ---- checkpoint hooks (engine.hpp:221-232) --------------------------------------------------
    def set_dump_callback(self, fn: Callable):
        self.dump_callback = fn

    def run_dump_callback(self):
        if self.dump_callback:
            self.dump_callback()

    @traced("engine.checkpoint_table")
    def checkpoint_table(self, table_id: int, client: Optional[KVClientTable] = None):
        """KVClientTable::CheckPoint (kv_client_table.hpp:136-154).  Collective on this backend."""
        from ..utils import checkpoint as ckpt
        ckpt.dump_table(self, self.tables[table_id], client)

    def restore_table(self, table_id: int) -> Dict[int, int]:
        """Model Restore (ssp_model.cpp:127-137): reload shards + iteration map from the last dump."""
        from ..utils import checkpoint as ckpt
        it_map = ckpt.restore_table(self, self.tables[table_id])
        self.iteration_map.update(it_map)
        return it_map

    def rollback(self) -> Dict[int, int]:
        """Global rollback to the last checkpoint (comm/mailbox.cpp:172-190): every table is
        restored; workers resume at the checkpointed iteration."""
        out: Dict[int, int] = {}
        for tid in sorted(self.tables):
            if self.tables[tid].hview is None:
                out = self.restore_table(tid)
        return out

    def set_iteration(self, worker_id: int, iteration: int):
        self.iteration_map[worker_id] = iteration

    def get_iteration(self, worker_id: int) -> int:
        return self.iteration_map.get(worker_id, self.default_iteration)

    # ---- convenience ----------------------------------------------------------------------------
    @classmethod
    def from_env(cls, device: Optional[int] = None) -> "Engine":
        """Build from torchrun-style env (RANK/LOCAL_RANK/WORLD_SIZE/MASTER_*), NCCL for the
        baseline collectives + gloo control plane."""
        import torch.distributed as dist
        world = int(os.environ.get("WORLD_SIZE", "1"))
        rank = int(os.environ.get("RANK", "0"))
        local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if device is None:
            device = local_rank
        numa = None
        if torch.cuda.is_available():
            torch.cuda.set_device(device)
            from ..utils.numa import bind_to_gpu_node
            numa = bind_to_gpu_node(device)
        if world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            dist.init_process_group(backend="nccl" if torch.cuda.is_available() else "gloo",
                                    rank=rank, world_size=world)
        nodes = [Node(i, "localhost", 0) for i in range(world)]
        eng = cls(nodes[rank], nodes, device=device)
        eng.numa = numa
        return eng
