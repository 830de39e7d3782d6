"""GPU numerics of the one-sided push / pull / clock kernels against plain PyTorch fp32.

One GPU, several shards on it (num_server_threads_per_node > 1): the routing, staging and
fold logic is exactly the multi-GPU code path with every "peer" pointer being local.
"""
import pytest
import torch

from minips_b200.parallel.types import MLTask, ModelType, Node, StorageType, WorkerAlloc

pytestmark = pytest.mark.gpu


@pytest.fixture()
def engine():
    from minips_b200.parallel.engine import Engine
    n = Node(0, "localhost", 0)
    e = Engine(n, [n], device=0)
    e.start_everything(num_server_threads_per_node=8)
    yield e
    e.stop_everything()


def _run(engine, tid, fn, workers=1):
    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(0, workers)])
    task.set_tables([tid])
    task.set_lambda(fn)
    engine.run(task)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("model", [ModelType.ASP, ModelType.SSP, ModelType.BSP])
def test_add_get_roundtrip(engine, dtype, model):
    num_dims = 100_003  # not a multiple of 8 or 4: exercises the ragged last shard
    tid = engine.create_table(num_dims=num_dims, model_type=model, storage_type=StorageType.Vector,
                              staleness=1, dtype=dtype)
    g = torch.Generator().manual_seed(1)
    keys = torch.sort(torch.randperm(num_dims, generator=g)[:20000])[0]
    vals = torch.randn(keys.numel(), generator=g, dtype=dtype)
    out = {}

    def fn(info):
        t = info.create_kv_client_table(tid)
        t.add(keys, vals)
        t.add(keys, vals)  # Add accumulates (server/vector_storage.hpp:35)
        t.clock()
        out["got"] = t.get(keys).cpu()
        out["all"] = t.get_all().cpu()

    _run(engine, tid, fn)
    ref = torch.zeros(num_dims, dtype=dtype)
    ref[keys] = 2 * vals
    torch.testing.assert_close(out["got"], 2 * vals)
    torch.testing.assert_close(out["all"], ref)


def test_bsp_add_invisible_until_clock(engine):
    """bsp_model_test.cpp:31-132: adds of clock t are not visible to gets of clock t."""
    tid = engine.create_table(num_dims=4096, model_type=ModelType.BSP)
    keys = torch.arange(0, 4096, 3)
    out = {}

    def fn(info):
        t = info.create_kv_client_table(tid)
        t.add(keys, torch.ones(keys.numel()))
        out["before"] = t.get(keys).cpu()
        t.clock()
        out["after"] = t.get(keys).cpu()

    _run(engine, tid, fn)
    assert torch.all(out["before"] == 0)
    assert torch.all(out["after"] == 1)


def test_out_of_range_keys_dropped(engine):
    """range_partition_manager.hpp:24-41: keys outside every range are silently dropped."""
    tid = engine.create_table(ranges=[(10 + 5 * i, 15 + 5 * i) for i in range(8)], model_type=ModelType.ASP)
    keys = torch.tensor([2, 10, 14, 30, 49, 50, 77])
    out = {}

    def fn(info):
        t = info.create_kv_client_table(tid)
        t.add(keys, torch.arange(1, 8, dtype=torch.float32))
        out["got"] = t.get(keys).cpu()

    _run(engine, tid, fn)
    assert out["got"].tolist() == [0, 2, 3, 4, 5, 0, 0]


def test_vals_per_key(engine):
    tid = engine.create_table(num_dims=1000, model_type=ModelType.ASP, vals_per_key=3)
    keys = torch.tensor([1, 500, 999])
    vals = torch.arange(9, dtype=torch.float32)
    out = {}

    def fn(info):
        t = info.create_kv_client_table(tid)
        t.add(keys, vals)
        out["got"] = t.get(keys).cpu()

    _run(engine, tid, fn)
    torch.testing.assert_close(out["got"], vals)


def test_dense_push_pull(engine):
    num_dims = 1_000_003
    tid = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
    delta = torch.randn(num_dims, device="cuda")
    delta[::7] = 0
    ref = delta.clone()
    out = {}

    def fn(info):
        t = info.create_kv_client_table(tid)
        d = delta.clone()
        t.add_all(d, zero_src=True)
        out["zeroed"] = bool((d == 0).all())
        t.clock()
        t.add_all(delta.clone())
        t.clock()
        out["all"] = t.get_all()

    _run(engine, tid, fn)
    assert out["zeroed"]
    torch.testing.assert_close(out["all"], 2 * ref)


def test_basic_example_semantics(engine):
    """apps/basic/basic_example.cpp:36-70: W workers x 100 x (Get, Add 0.5, Clock) under SSP."""
    W, iters = 4, 25
    tid = engine.create_table(ranges=[(125 * i, 125 * (i + 1)) for i in range(8)], model_type=ModelType.SSP,
                              storage_type=StorageType.Map, staleness=1)
    keys = torch.arange(1000)

    def fn(info):
        t = info.create_kv_client_table(tid)
        for _ in range(iters):
            t.get(keys)
            t.add(keys, torch.full((1000,), 0.5))
            t.clock()

    _run(engine, tid, fn, workers=W)
    full = engine.tables[tid].gather_full().cpu()
    assert torch.all(full == 0.5 * iters * W)


def test_multi_worker_bsp_lockstep(engine):
    """Two workers on one device under BSP: every get sees exactly the folded adds of all
    workers of earlier clocks."""
    tid = engine.create_table(num_dims=512, model_type=ModelType.BSP)
    keys = torch.arange(512)
    seen = {0: [], 1: []}

    def fn(info):
        t = info.create_kv_client_table(tid)
        for _ in range(5):
            seen[info.worker_id].append(t.get(keys))
            t.add(keys, torch.ones(512))
            t.clock()

    _run(engine, tid, fn, workers=2)
    for w in (0, 1):
        for it, v in enumerate(seen[w]):
            assert torch.all(v.cpu() == 2.0 * it), (w, it, v[:4])


def test_device_barrier_single_rank(engine):
    tid = engine.create_table(num_dims=64, model_type=ModelType.ASP)

    def fn(info):
        for _ in range(3):
            engine.device_barrier(tid)
        torch.cuda.synchronize()

    _run(engine, tid, fn)
    assert engine.tables[tid].barrier_epoch == 3


def test_straggle_and_phase_timer():
    """mps_straggle holds the stream for the requested time; PhaseTimer attributes it to its phase."""
    import ctypes as C
    from minips_b200.ops import _lib
    from minips_b200.utils.tracing import PhaseTimer
    lib = _lib.load()
    pt = PhaseTimer()
    x = torch.zeros(1 << 20, device="cuda")
    for _ in range(3):
        with pt.phase("straggle"):
            _lib.check(lib.mps_straggle(C.c_longlong(2_000_000), _lib.stream_ptr()), "straggle")
        with pt.phase("work"):
            x.add_(1.0)
    s = pt.summary("cuda", reduce_over_ranks=False)
    assert s["straggle"]["calls"] == 3 and s["work"]["calls"] == 3
    assert 6.0 <= s["straggle"]["ms"] < 30.0, s
    assert s["work"]["ms"] < s["straggle"]["ms"]
