"""Failure detection and recovery for the device backend.

The reference's flow (master/heartbeat_check_thread.cpp:15-41, master/master_thread.cpp:19-43,
comm/mailbox.cpp:172-190, apps/logistic_regression/lr_example.cpp:265-288): every node sends heartbeats
to the master; a node silent for 3 intervals is declared failed and relaunched with
`use_weight_file`; the relaunched process's first heartbeat makes the master broadcast kRollBack; every
node reloads the last checkpoint and continues from the checkpointed iteration.

Here the control plane is the SAME C++ code (csrc/core/master.cpp: Master, MasterThread,
HeartBeatCheckThread; ControlAgent = mailbox + heartbeat of a rank without server / worker threads), bound
through `minips_b200.core`.  What differs is what "roll back" has to do on this backend: the shards of a
dead rank were its HBM, and the survivors' peer mappings and process group died with it.  So a rollback
is a new GENERATION: every rank — survivors and the relaunched one — tears its engine down, meets the
others again on a fresh process group, re-creates the tables, restores shards, clocks and iteration
from the checkpoint and resumes there.  The rendezvous store lives in the master process; the master
bumps `generation` right before it broadcasts kRollBack, so whoever receives the message finds it.

    master:   python -m minips_b200.apps.master --config_file=... --heartbeat_interval=1 --relaunch_cmd=...
    ranks:    python -m minips_b200.apps.lr_example --config_file=... --my_id=N --heartbeat_interval=1 ...

Elastic scale-out (reference: lr_example.cpp:396-409, driver/engine.hpp:81-114, master/master_thread.cpp:
40-43, comm/mailbox.cpp:197-219) is the same mechanism with a changed membership: a NEW process started
with --scale --my_id=K --scale_name=host --scale_port=P writes its address into the scale file, starts
its control agent and sends kScale; the master adds it to its node list, publishes the membership of the
next generation, bumps `generation` and broadcasts kScaleRollback.  Every rank — old and new — meets in
that generation with world + 1, the tables are partitioned over all of them (the new GPU's HBM takes its
share: the restore re-shards, utils/checkpoint.py) and training resumes at the checkpointed iteration.

    new rank: python -m minips_b200.apps.lr_example --config_file=... --scale --my_id=K --scale_name=host \
                  --scale_port=P --scale_file=/shared/scale --heartbeat_interval=1 ...
"""
from __future__ import annotations

import os
import time
from datetime import timedelta
from typing import Callable, List, Optional


class RollbackSignal(Exception):
    """Raised inside a generation when the master has ordered a rollback."""


def split_master(nodes, heartbeat_interval: int):
    """SelectMaster (base/node_utils.cpp:19-32): with heartbeats on, the node with id 1 is the master and
    leaves the worker list."""
    from .. import core
    master, rest = core.select_master(list(nodes), int(heartbeat_interval))
    return master, rest


class FaultTolerantJob:
    """One rank's side of the protocol.  `run(body)` calls body(job) once per generation until one
    generation completes; `job.check()` inside the body raises RollbackSignal as soon as the master has
    ordered a rollback."""

    def __init__(self, nodes, my_id: int, heartbeat_interval: int, store_port: int, relaunched: bool = False,
                 pg_timeout_s: float = 60.0, rollback_wait_s: Optional[float] = None, backend: str = "gloo",
                 scale_node=None, scale_file: str = ""):
        from .. import core
        self.core = core
        self.master, self.rank_nodes = split_master(nodes, heartbeat_interval)
        assert self.master.is_master, "fault tolerance needs heartbeat_interval > 0 and a node with id 1"
        self.scale_node = scale_node          # this process joins a running job as a new node
        self.scale_file = scale_file
        if scale_node is not None:
            assert scale_node.id == my_id and my_id not in [n.id for n in nodes], "a scale node needs a fresh id"
            self.rank_nodes = list(self.rank_nodes) + [scale_node]
        self.known = {n.id: n for n in self.rank_nodes}
        ids = [n.id for n in self.rank_nodes]
        assert my_id in ids, f"node {my_id} is not a worker node of this cluster"
        self.my_id = my_id
        self.rank = ids.index(my_id)          # both are re-read from the store at every generation
        self.world = len(ids)
        self.me = self.rank_nodes[self.rank]
        self.interval = int(heartbeat_interval)
        self.store_port = int(store_port)
        self.relaunched = relaunched
        self.pg_timeout_s = pg_timeout_s
        self.rollback_wait_s = rollback_wait_s if rollback_wait_s is not None else 6 * self.interval + 90
        self.backend = backend
        self.agent = None
        self.store = None
        self.generation = 0
        self.generations_run = 0

    # ---- control plane ----------------------------------------------------------------------------
    def start(self):
        import torch.distributed as dist
        core = self.core
        core.context_set("heartbeat_interval", str(self.interval))
        core.context_set("my_id", str(self.my_id))
        if self.relaunched:
            core.check_fault_tolerance(3)                     # "failed process restarted" (lr_example.cpp:120)
        self.agent = core.ControlAgent(self.me, list(self.rank_nodes), self.master)
        self.agent.start()
        host = "127.0.0.1" if self.master.hostname == "localhost" else self.master.hostname
        self.store = dist.TCPStore(host, self.store_port, is_master=False, timeout=timedelta(seconds=120),
                                   wait_for_workers=False)
        if self.scale_node is not None:
            # lr_example.cpp:396-409: announce the address (scale file + store), then kScale; the master's
            # answer is kScaleRollback and a new generation whose membership includes this node
            before = int(self.store.add("generation", 0))
            self.store.set(f"node/{self.my_id}", f"{self.me.hostname}:{self.me.port}")
            if self.scale_file:
                core.dump_scale_file(self.me, self.scale_file)
            self.agent.scale()
            end = time.time() + self.rollback_wait_s
            while time.time() < end and int(self.store.add("generation", 0)) == before:
                time.sleep(0.05)
            assert int(self.store.add("generation", 0)) > before, "the master did not answer kScale"
            self._await_rollback(5.0)         # the kScaleRollback addressed to this node: consumed here
            self.resume = True
        elif self.relaunched:
            # my first heartbeat makes the master order the rollback; the generation to join is the one it
            # publishes with that order
            self._await_rollback(self.rollback_wait_s)

    def stop(self):
        if self.agent is not None:
            self.agent.stop(True)
            self.agent = None

    def check(self):
        if self.agent is not None and self.agent.need_rollback():
            if int(self.store.add("generation", 0)) > self.generation:
                raise RollbackSignal()
            self.agent.clear_rollback()       # an order for the generation this process already runs in

    def sleep(self, seconds: float):
        """Host-side wait that stays responsive to a rollback order."""
        end = time.time() + seconds
        while time.time() < end:
            self.check()
            time.sleep(0.05)

    def _await_rollback(self, seconds: float) -> bool:
        end = time.time() + seconds
        while time.time() < end:
            if self.agent.need_rollback():
                self.agent.clear_rollback()
                return True
            time.sleep(0.05)
        return False

    # ---- generations --------------------------------------------------------------------------------
    def _members(self, generation: int):
        """Node ids of a generation, in rank order (published by the master before it bumps `generation`)."""
        from .. import core
        key = f"members/{generation}"
        self.store.wait([key])
        ids = [int(x) for x in self.store.get(key).decode().split(",") if x]
        for i in ids:
            if i not in self.known:           # a node that scaled in after this process started
                host, port = self.store.get(f"node/{i}").decode().rsplit(":", 1)
                self.known[i] = core.Node(i, host, int(port))
        return ids

    def _join(self):
        import torch.distributed as dist
        self.generation = int(self.store.add("generation", 0))
        ids = self._members(self.generation)
        assert self.my_id in ids, f"node {self.my_id} is not a member of generation {self.generation}: {ids}"
        self.rank_nodes = [self.known[i] for i in ids]
        self.rank, self.world = ids.index(self.my_id), len(ids)
        prefix = dist.PrefixStore(f"gen{self.generation}/", self.store)
        dist.init_process_group(self.backend, store=prefix, rank=self.rank, world_size=self.world,
                                timeout=timedelta(seconds=self.pg_timeout_s))

    @staticmethod
    def _leave():
        import torch.distributed as dist
        try:
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001 - the group is already broken
            pass

    def run(self, body: Callable[["FaultTolerantJob"], None]):
        """body(job) builds the engine, trains (resuming from the checkpoint when job.resume) and tears the
        engine down.  Any failure inside a generation — a device-side wait that timed out on a dead peer,
        a broken process group, an explicit RollbackSignal — is held until the master's order arrives; if
        none does within rollback_wait_s the failure is re-raised."""
        self.resume = getattr(self, "resume", False) or self.relaunched
        while True:
            if self.resume:
                # torch's symmetric-memory allocator keeps per-group state of the process group that died
                # (a rendezvous on the new group waits for its time-out): generations after a rollback map
                # their buffers through CUDA IPC, peer loads / stores instead of NVLS
                os.environ["MINIPS_HEAP"] = "ipc"
            self._join()
            self.generations_run += 1
            failure: Optional[BaseException] = None
            try:
                if self.resume:
                    import torch.distributed as dist
                    dist.barrier()
                    self.core.check_fault_tolerance(5)        # "all nodes recovered" (lr_example.cpp:266)
                body(self)
            except BaseException as e:  # noqa: BLE001
                failure = e
            self._leave()
            if failure is None:
                return
            if isinstance(failure, KeyboardInterrupt):
                raise failure
            print(f"[node {self.my_id}] generation {self.generation} ended with {type(failure).__name__}: "
                  f"{str(failure)[:200]}; waiting for the master's rollback order", flush=True)
            if not self._await_rollback(self.rollback_wait_s):
                raise failure
            self.resume = True


class MasterProcess:
    """The master node (id 1): C++ Master + heartbeat check thread, plus the rendezvous store of the
    generations (reference master/master.hpp:14-87)."""

    def __init__(self, nodes, heartbeat_interval: int, relaunch_cmd: str, store_port: int, scale_file: str = ""):
        from .. import core
        import torch.distributed as dist
        self.core = core
        core.context_set("heartbeat_interval", str(int(heartbeat_interval)))
        core.context_set("relaunch_cmd", relaunch_cmd)
        self.master_node, self.rank_nodes = split_master(nodes, heartbeat_interval)
        assert self.master_node.is_master
        self.store = dist.TCPStore("0.0.0.0", int(store_port), is_master=True, timeout=timedelta(seconds=120),
                                   wait_for_workers=False)
        self.store.add("generation", 0)
        self.members: List[int] = [int(n.id) for n in self.rank_nodes]
        self.generation = 0
        self.store.set("members/0", ",".join(str(i) for i in self.members))
        if scale_file:
            core.context_set("scale_file", scale_file)
        self.master = core.Master(self.master_node, list(self.rank_nodes))
        self.rollbacks: List[int] = []
        self.scaled: List[int] = []

        def next_generation():
            # membership first, then the counter: whoever sees the new generation finds its members
            self.generation += 1
            self.store.set(f"members/{self.generation}", ",".join(str(i) for i in self.members))
            self.store.add("generation", 1)

        def on_rollback(failed_id: int):
            self.rollbacks.append(int(failed_id))
            next_generation()

        def on_scale(node_id: int):
            # master/master_thread.cpp:40-43: kScale -> ScaleRollBack; the new node joins the next generation
            self.scaled.append(int(node_id))
            if int(node_id) not in self.members:
                self.members.append(int(node_id))
            next_generation()

        self.master.set_rollback_hook(on_rollback)
        self.master.set_scale_hook(on_scale)

    def serve(self):
        """Blocks until every node has said goodbye (kQuitHeartBeat)."""
        self.master.start_master()
        self.master.stop_master()
