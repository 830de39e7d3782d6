"""Ports of the reference's driver-level unit tests (Python driver vocabulary).

driver/simple_id_mapper_test.cpp:22-75, driver/worker_spec_test.cpp:23-86, ml_task basics.
"""
import pytest

from minips_b200.parallel.types import (MLTask, Node, SimpleIdMapper, WorkerAlloc, WorkerSpec, even_ranges)


def _mapper():
    nodes = [Node(0, "localhost", 12351), Node(1, "localhost", 12352), Node(3, "localhost", 12353)]
    m = SimpleIdMapper(nodes[0], nodes)
    return m


def test_id_mapper_init_layout():
    m = _mapper()
    m.init(2)
    assert m.get_node_id_for_thread(0) == 0
    assert m.get_node_id_for_thread(1000) == 1
    assert m.get_node_id_for_thread(3001) == 3
    assert m.get_server_threads_for_id(0) == [0, 1]
    assert m.get_server_threads_for_id(1) == [1000, 1001]
    assert m.get_server_threads_for_id(3) == [3000, 3001]
    assert m.get_worker_helper_threads_for_id(0) == [50]
    assert m.get_worker_helper_threads_for_id(1) == [1050]
    assert m.get_worker_helper_threads_for_id(3) == [3050]
    assert m.get_all_server_threads() == [0, 1, 1000, 1001, 3000, 3001]


def test_id_mapper_allocate_deallocate():
    m = _mapper()
    m.init(1)
    assert m.allocate_worker_thread(0) == 100
    assert m.allocate_worker_thread(0) == 101
    assert m.allocate_worker_thread(1) == 1100
    assert m.allocate_worker_thread(1) == 1101
    assert m.allocate_worker_thread(1) == 1102
    assert m.get_worker_threads_for_id(1) == [1100, 1101, 1102]
    m.deallocate_worker_thread(1, 1101)
    assert m.get_worker_threads_for_id(1) == [1100, 1102]
    assert m.allocate_worker_thread(1) == 1101  # first free id is reused


def test_worker_spec_maps():
    spec = WorkerSpec([WorkerAlloc(0, 3), WorkerAlloc(1, 2)])
    assert spec.num_workers == 5
    assert spec.has_local_workers(0) and spec.has_local_workers(1) and not spec.has_local_workers(2)
    assert spec.get_local_workers(0) == [0, 1, 2]
    assert spec.get_local_workers(1) == [3, 4]
    assert [spec.worker_to_node[w] for w in range(5)] == [0, 0, 0, 1, 1]
    spec.insert_worker_id_thread_id(0, 100)
    spec.insert_worker_id_thread_id(1, 101)
    spec.insert_worker_id_thread_id(3, 1100)
    assert spec.get_all_thread_ids() == [100, 101, 1100]
    assert spec.get_local_threads(0) == [100, 101]
    assert spec.get_local_threads(1) == [1100]
    assert spec.worker_to_thread[3] == 1100 and spec.thread_to_worker[101] == 1
    with pytest.raises(AssertionError):
        spec.insert_worker_id_thread_id(0, 102)


def test_ml_task_setup():
    t = MLTask()
    assert not t.is_setup()
    t.set_lambda(lambda info: None)
    t.set_worker_alloc([(0, 2), WorkerAlloc(1, 1)])
    assert not t.is_setup()
    t.set_tables([0])
    assert t.is_setup()
    assert t.get_worker_alloc()[0].num_workers == 2


def test_even_ranges_last_takes_remainder():
    # Engine::getRanges, driver/engine.hpp:158-172
    assert even_ranges(10, 3) == [(0, 3), (3, 6), (6, 10)]
    assert even_ranges(8, 8)[-1] == (7, 8)


def test_tracing_helpers_are_noops_without_cuda_or_flag():
    from minips_b200.utils.tracing import PhaseTimer, nvtx_range, traced

    @traced("x")
    def f(a):
        return a + 1

    with nvtx_range("outer"):
        assert f(1) == 2
    pt = PhaseTimer(enabled=False)
    with pt.phase("p"):
        pass
    assert pt.summary() == {}


def test_packed_ids_bit_stream_round_trips():
    """Loader's compact id stream: ids at 16 / 20 / 24 bits, little-endian, decodable at any start offset."""
    import torch
    from minips_b200.utils.loader import PackedCsrHost
    for bits, top in ((16, 60_000), (20, 1_000_000), (24, 10_000_000)):
        col = torch.randint(0, top, (1001,), dtype=torch.int32)
        stream = int.from_bytes(PackedCsrHost.pack_ids(col, bits).numpy().tobytes(), "little")
        got = [(stream >> (i * bits)) & ((1 << bits) - 1) for i in range(col.numel())]
        assert got == col.tolist(), bits
