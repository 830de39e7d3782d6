"""The C++ host runtime through its pybind11 bindings: reference driver API from Python on CPU."""
import os
import threading

import pytest

core = pytest.importorskip("minips_b200.core")


def _port():
    return 29000 + os.getpid() % 1500


def test_partition_manager_and_tracker():
    pm = core.RangePartitionManager([0, 1, 2], [(2, 4), (4, 7), (7, 10)])
    assert pm.slice_keys([2, 8, 9]) == [(0, [2]), (2, [8, 9])]
    kv = pm.slice_kvs([2, 5, 9], [0.2, 0.5, 0.9])
    assert [k[0] for k in kv] == [0, 1, 2] and kv[1][2] == [0.5]
    t = core.ProgressTracker()
    t.init([1, 3])
    assert [t.advance_and_get_changed_min_clock(i) for i in (1, 3, 1, 1, 3)] == [-1, 1, -1, -1, 2]
    assert core.parse_libsvm("-1 3:1 11:0.5") == ([(2, 1.0), (10, 0.5)], -1.0)
    assert core.parse_libsvm("# comment") is None


def test_engine_bsp_from_python():
    node = core.Node(0, "localhost", _port())
    e = core.Engine(node, [node])
    e.start_everything(2)
    tid = e.create_table([(0, 50), (50, 100)], core.ModelType.BSP, core.StorageType.Vector, 0)
    e.barrier()
    seen = {}
    lock = threading.Lock()

    def fn(info):
        t = info.create_kv_client_table(tid)
        keys = list(range(0, 100, 7))
        mine = []
        for _ in range(4):
            mine.append(t.get(keys))
            t.add(keys, [1.0] * len(keys))
            t.clock()
        with lock:
            seen[info.worker_id] = mine

    task = core.MLTask()
    task.set_lambda(fn)
    task.set_worker_alloc([core.WorkerAlloc(0, 3)])
    task.set_tables([tid])
    assert task.is_setup()
    e.run(task)
    e.stop_everything()
    assert len(seen) == 3
    for vals in seen.values():
        for it, v in enumerate(vals):
            assert v == [3.0 * it] * len(v)  # BSP: all three workers' adds of earlier clocks, none of this one
