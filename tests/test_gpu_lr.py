"""Fused sparse-LR kernels against a plain PyTorch fp32 oracle of the same iteration."""
import pytest
import torch

from minips_b200.parallel.types import MLTask, ModelType, Node, WorkerAlloc

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


def _oracle(data, num_dims, starts, batch, alpha):
    from minips_b200.models.lr import oracle_step
    w = torch.zeros(num_dims, device="cuda")
    stats = []
    for s in starts:
        d, loss, correct = oracle_step(w, data, int(s), batch, alpha)
        w = w + d
        stats.append((loss, correct))
    return w, stats


def _oracle64(data, num_dims, starts, batch, alpha):
    """The same iteration in fp64 PyTorch (the reference's Val = double)."""
    from minips_b200.models.lr import CsrBatchSource, oracle_step
    d64 = CsrBatchSource(data.row_ptr, data.col, data.val.double(), data.label.double())
    w = torch.zeros(num_dims, device="cuda", dtype=torch.float64)
    stats = []
    for s in starts:
        d, loss, correct = oracle_step(w, d64, int(s), batch, alpha)
        w = w + d
        stats.append((loss, correct))
    return w, stats


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("model,protocol", [(ModelType.BSP, "xchg"), (ModelType.BSP, "pushpull"),
                                            (ModelType.SSP, "auto"), (ModelType.ASP, "auto")])
def test_dense_train_matches_oracle(engine, model, protocol, dtype):
    from minips_b200.models.lr import SparseLRTrainer, make_synthetic
    num_dims, n_rows, batch, iters, alpha = 50_001, 20_000, 4096, 6, 0.05
    data = make_synthetic(n_rows, num_dims, 40, seed=3)
    tid = engine.create_table(num_dims=num_dims, model_type=model, staleness=2, dtype=dtype)
    starts = torch.tensor([0, 4096, 17_000, 19_999, 5, 12_345], dtype=torch.int64, device="cuda")
    out = {}

    def fn(info):
        tr = SparseLRTrainer(info.create_kv_client_table(tid), data, alpha)
        # two launches: the second one starts from the state the first one left (replica, parity of
        # the accumulators, clocks)
        l0 = tr.train_dense(starts[:4], batch, protocol=protocol)
        l1 = tr.train_dense(starts[4:], batch, protocol=protocol)
        out["loss"] = torch.cat([l0, l1])
        out["w"] = tr.client.get_all()
        if tr._gacc is not None:
            out["gacc_clean"] = bool((tr._gacc == 0).all())
        else:
            out["replica_is_w"] = bool(torch.equal(engine.tables[tid].xchg.replica, out["w"]))
        out["eval"] = tr.evaluate(out["w"])

    _run(engine, tid, fn)
    if dtype == torch.float64:
        w_ref, stats = _oracle64(data, num_dims, starts.tolist(), batch, alpha)
        torch.testing.assert_close(out["w"], w_ref, rtol=1e-9, atol=1e-12)
    else:
        w_ref, stats = _oracle(data, num_dims, starts.tolist(), batch, alpha)
        torch.testing.assert_close(out["w"], w_ref, rtol=2e-4, atol=2e-5)
    loss = out["loss"].cpu()
    for i, (l, c) in enumerate(stats):
        assert abs(loss[i, 0].item() - l) <= 2e-3 * abs(l) + 1e-2
        assert abs(loss[i, 1].item() - c) <= 2
    assert out.get("gacc_clean", True) and out.get("replica_is_w", True)
    mean_loss, acc = out["eval"]
    assert 0.0 < mean_loss < 0.75 and 0.5 <= acc <= 1.0


def test_xchg_replica_refreshed_after_generic_add(engine):
    """A generic Add between two fused launches makes the kept replica stale; the trainer notices its
    own writes and the kernel re-reads the shards."""
    from minips_b200.models.lr import SparseLRTrainer, make_synthetic, oracle_step
    num_dims, n_rows, batch, alpha = 30_000, 9000, 2048, 0.05
    data = make_synthetic(n_rows, num_dims, 40, seed=13)
    tid = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
    bump = torch.linspace(-0.5, 0.5, num_dims, device="cuda")
    out = {}

    def fn(info):
        c = info.create_kv_client_table(tid)
        tr = SparseLRTrainer(c, data, alpha)
        tr.train_dense(torch.tensor([0], device="cuda"), batch)
        c.add(torch.arange(num_dims), bump)
        c.clock()
        tr.train_dense(torch.tensor([4000], device="cuda"), batch)
        out["w"] = c.get_all()

    _run(engine, tid, fn)
    w = torch.zeros(num_dims, device="cuda")
    w = w + oracle_step(w, data, 0, batch, alpha)[0] + bump
    w = w + oracle_step(w, data, 4000, batch, alpha)[0]
    torch.testing.assert_close(out["w"], w, rtol=2e-4, atol=2e-5)


def test_out_of_range_feature_ids_are_dropped(engine):
    """Feature ids >= num_dims (or the 0xFFFFFFFF a libsvm index of 0 parses to) must neither be read
    nor added anywhere; the reference's Slice drops them (range_partition_manager.hpp:24-41)."""
    from minips_b200.models.lr import CsrBatchSource, SparseLRTrainer, make_synthetic
    num_dims, n_rows, batch, alpha = 10_000, 2048, 2048, 0.05
    data = make_synthetic(n_rows, num_dims, 20, seed=17)
    col = data.col.clone()
    col[::7] = num_dims + 5
    col[3::11] = -1                                     # 0xFFFFFFFF
    bad = CsrBatchSource(data.row_ptr, col, data.val, data.label)
    keep = (col >= 0) & (col < num_dims)
    clean = CsrBatchSource(data.row_ptr, torch.where(keep, col, torch.zeros_like(col)),
                           torch.where(keep, data.val, torch.zeros_like(data.val)), data.label)
    tid = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
    tid2 = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
    out = {}
    starts = torch.zeros(2, dtype=torch.int64, device="cuda")

    def fn(info):
        a = SparseLRTrainer(info.create_kv_client_table(tid), bad, alpha)
        a.train_dense(starts, batch)
        out["bad"] = a.client.get_all()
        out["eval"] = a.evaluate(out["bad"])
        b = SparseLRTrainer(info.create_kv_client_table(tid2), bad, alpha)
        b.train_dense(starts, batch, protocol="pushpull")
        out["bad_pp"] = b.client.get_all()

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(0, 1)])
    task.set_tables([tid, tid2])
    task.set_lambda(fn)
    engine.run(task)
    w_ref, _ = _oracle(clean, num_dims, [0, 0], batch, alpha)
    torch.testing.assert_close(out["bad"], w_ref, rtol=3e-4, atol=3e-5)
    torch.testing.assert_close(out["bad_pp"], w_ref, rtol=3e-4, atol=3e-5)
    assert 0.0 < out["eval"][0] < 1.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("model", [ModelType.BSP, ModelType.ASP])
def test_sparse_step_matches_oracle(engine, model, dtype):
    from minips_b200.models.lr import SparseLRTrainer, make_synthetic
    num_dims, n_rows, batch, alpha = 2_000_003, 6000, 1024, 0.1
    data = make_synthetic(n_rows, num_dims, 40, seed=5)
    tid = engine.create_table(num_dims=num_dims, model_type=model, dtype=dtype)
    starts = [0, 1024, 5500]
    out = {}

    def fn(info):
        tr = SparseLRTrainer(info.create_kv_client_table(tid), data, alpha)
        losses = [tr.step_sparse(s, batch) for s in starts]
        out["w"] = tr.client.get_all()
        out["loss"] = torch.stack(losses).cpu()

    _run(engine, tid, fn)
    if dtype == torch.float64:
        w_ref, stats = _oracle64(data, num_dims, starts, batch, alpha)
    else:
        w_ref, stats = _oracle(data, num_dims, starts, batch, alpha)
    if model == ModelType.BSP and dtype == torch.float64:
        torch.testing.assert_close(out["w"], w_ref, rtol=1e-9, atol=1e-12)
    elif model == ModelType.BSP:
        torch.testing.assert_close(out["w"], w_ref, rtol=2e-4, atol=2e-5)
    else:
        # In-place adds (SSP/ASP) are visible to rows of the same batch that are still being
        # processed — asynchronous by design (Hogwild within the step).  The handful of keys
        # shared by concurrently processed rows may differ from the synchronous oracle by a
        # second-order term; everything else matches tightly.
        bad = ~torch.isclose(out["w"], w_ref, rtol=2e-4, atol=2e-5)
        assert bad.float().mean().item() < 1e-3
        torch.testing.assert_close(out["w"], w_ref, rtol=0, atol=5e-3)
    for i, (l, c) in enumerate(stats):
        assert abs(out["loss"][i, 0].item() - l) <= 2e-3 * abs(l) + 1e-2


def test_long_rows(engine):
    """Rows with more than 64 non-zeros take the re-read path of both kernels."""
    from minips_b200.models.lr import SparseLRTrainer, make_synthetic
    num_dims, n_rows, batch, alpha = 10_000, 512, 512, 0.01
    data = make_synthetic(n_rows, num_dims, 150, seed=7)
    tid_d = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
    tid_s = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP)
    out = {}
    starts = torch.zeros(2, dtype=torch.int64, device="cuda")

    def fn(info):
        a = SparseLRTrainer(info.create_kv_client_table(tid_d), data, alpha)
        a.train_dense(starts, batch)
        out["dense"] = a.client.get_all()
        b = SparseLRTrainer(info.create_kv_client_table(tid_s), data, alpha)
        b.step_sparse(0, batch)
        b.step_sparse(0, batch)
        out["sparse"] = b.client.get_all()

    task = MLTask()
    task.set_worker_alloc([WorkerAlloc(0, 1)])
    task.set_tables([tid_d, tid_s])
    task.set_lambda(fn)
    engine.run(task)
    w_ref, _ = _oracle(data, num_dims, [0, 0], batch, alpha)
    torch.testing.assert_close(out["dense"], w_ref, rtol=3e-4, atol=3e-5)
    torch.testing.assert_close(out["sparse"], w_ref, rtol=3e-4, atol=3e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_stream_reads_packed_batches_in_place(engine, dtype):
    """End-to-end steps: batches arrive from pinned host memory in the loader's compact form and the
    training kernel decodes them itself (row offsets from the per-8-rows table, bit-packed ids, u8
    labels) — one launch per step.  Same result as the plain-CSR oracle."""
    from minips_b200.models.lr import SparseLRTrainer, make_synthetic
    from minips_b200.utils.loader import AsyncBatchLoader, PackedCsrHost
    num_dims, n_rows, batch, alpha = 70_001, 9000, 2048, 0.05
    g = torch.Generator(device="cuda").manual_seed(5)
    # ragged rows (0..60 non-zeros) so that the offset arithmetic is exercised
    lens = torch.randint(0, 61, (n_rows,), device="cuda", generator=g)
    rp = torch.zeros(n_rows + 1, dtype=torch.int64, device="cuda")
    rp[1:] = torch.cumsum(lens, 0)
    nnz = int(rp[-1])
    from minips_b200.models.lr import CsrBatchSource
    col = torch.randint(0, num_dims, (nnz,), device="cuda", generator=g, dtype=torch.int64).to(torch.int32)
    val = torch.randn(nnz, device="cuda", generator=g)
    lab = (torch.rand(n_rows, device="cuda", generator=g) < 0.5).float()
    data = CsrBatchSource(rp, col, val, lab)
    tid = engine.create_table(num_dims=num_dims, model_type=ModelType.BSP, dtype=dtype)
    starts = [0, 3001, 6951]
    out = {}

    def fn(info):
        tr = SparseLRTrainer(info.create_kv_client_table(tid), data, alpha)
        packed = PackedCsrHost.from_csr(data.to("cpu"))
        assert packed is not None and packed.id_bits == 20
        loader = AsyncBatchLoader(packed, batch, "cuda", depth=2, in_place=True)
        assert loader.in_place and loader.launches_per_batch == 0
        out["loss"] = tr.train_stream(loader, starts)
        torch.cuda.synchronize()
        out["w"] = tr.client.get_all()

    _run(engine, tid, fn)
    if dtype == torch.float64:
        w_ref, stats = _oracle64(data, num_dims, starts, batch, alpha)
        torch.testing.assert_close(out["w"], w_ref, rtol=1e-9, atol=1e-12)
    else:
        w_ref, stats = _oracle(data, num_dims, starts, batch, alpha)
        torch.testing.assert_close(out["w"], w_ref, rtol=2e-4, atol=2e-5)
    for i, (l, c) in enumerate(stats):
        assert abs(out["loss"][i, 0].item() - l) <= 2e-3 * abs(l) + 1e-2
