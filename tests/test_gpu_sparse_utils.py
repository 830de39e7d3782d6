"""Device sparse helpers vs their host definitions: batch unique keys (std::set semantics of
lib/batch_data_sampler.cpp:50-71), libsvm text -> CSR (lib/parser.hpp:17-47) and the
hash-backed MapStorage (server/map_storage.hpp:13-26)."""
import pytest
import torch

from minips_b200.parallel.types import MLTask, ModelType, Node, StorageType, WorkerAlloc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("num_dims,rows,batch,start", [(1000, 500, 64, 10), (1_000_003, 20_000, 4096, 777),
                                                       (5_000_000, 3000, 3000, 0)])
def test_batch_unique_keys(num_dims, rows, batch, start):
    from minips_b200.models.lr import make_synthetic
    from minips_b200.ops.sparse import batch_unique_keys
    data = make_synthetic(rows, num_dims, 40, seed=2)
    keys, slots = batch_unique_keys(data, start, batch, num_dims, want_slots=True)
    lo, hi = int(data.row_ptr[start]), int(data.row_ptr[start + batch])
    cols = data.col[lo:hi].to(torch.int64)
    ref = torch.unique(cols)  # sorted ascending, like std::set iteration order
    assert keys.numel() == ref.numel()
    assert torch.equal(keys.to(torch.int64), ref)
    # every non-zero maps to the slot holding its own key
    assert torch.equal(keys.to(torch.int64)[slots.long()], cols)


def test_batch_unique_keys_wraps_like_the_sampler():
    from minips_b200.models.lr import make_synthetic
    from minips_b200.ops.sparse import batch_unique_keys
    data = make_synthetic(100, 5000, 8, seed=4)
    keys, _ = batch_unique_keys(data, 90, 20, 5000)  # rows 90..99 then 0..9
    cols = torch.cat([data.col[90 * 8:], data.col[:10 * 8]]).to(torch.int64)
    assert torch.equal(keys.to(torch.int64), torch.unique(cols))


def _py_parse(text):
    rows = []
    for line in text.split("\n"):
        s = line.strip()
        if not s or s.startswith("#"):
            continue
        tok = s.split()
        y = float(tok[0])
        feats = []
        for t in tok[1:]:
            if t.startswith("#"):
                break
            i, v = t.split(":")
            feats.append((int(i) - 1, float(v)))
        rows.append((max(y, 0.0) if y < 0 else y, feats))
    return rows


def test_libsvm_to_csr_matches_host_parser(tmp_path):
    from minips_b200.ops.sparse import libsvm_to_csr, load_libsvm_file, write_synthetic_libsvm
    text = ("-1 3:1 11:0.5\t40:2e-1\n"
            "\n"
            "# a comment line\n"
            "+1 1:-1.25e+1 7:3.\r\n"
            "1\n"
            "  0 5:1 6:2 7:3 8:4   \n"
            "-1 123:4.5")  # no trailing newline
    csr = libsvm_to_csr(torch.tensor(list(text.encode()), dtype=torch.uint8, device="cuda"))
    ref = _py_parse(text)
    assert csr.n_rows == len(ref) == 5
    rp = csr.row_ptr.tolist()
    for r, (y, feats) in enumerate(ref):
        assert csr.label[r].item() == pytest.approx(y)
        assert rp[r + 1] - rp[r] == len(feats)
        for j, (i, v) in enumerate(feats):
            assert csr.col[rp[r] + j].item() == i
            assert csr.val[rp[r] + j].item() == pytest.approx(v, rel=1e-6)
    # a real file: BASELINE config-1 shape (10k x 1k)
    path = write_synthetic_libsvm(str(tmp_path / "a.libsvm"), 10_000, 1000, 10, seed=3)
    big = load_libsvm_file(path)
    ref = _py_parse(open(path).read())
    assert big.n_rows == 10_000 and big.nnz == 100_000
    assert int(big.col.max()) < 1000 and int(big.col.min()) >= 0
    for r in (0, 1234, 9999):
        lo, hi = int(big.row_ptr[r]), int(big.row_ptr[r + 1])
        assert [int(c) for c in big.col[lo:hi]] == [i for i, _ in ref[r][1]]
        torch.testing.assert_close(big.val[lo:hi].cpu(), torch.tensor([v for _, v in ref[r][1]]), rtol=1e-5, atol=1e-7)
        assert big.label[r].item() == ref[r][0]


def test_hash_backed_map_storage():
    from minips_b200.parallel.engine import Engine
    node = Node(0, "localhost", 0)
    engine = Engine(node, [node], device=0)
    engine.start_everything(num_server_threads_per_node=4)
    try:
        # 4 G-key space, only the touched keys cost memory
        ranges = [(i * (1 << 30), (i + 1) * (1 << 30)) for i in range(3)] + [(3 << 30, (1 << 32) - 1)]
        tid = engine.create_table(ranges=ranges, model_type=ModelType.SSP, storage_type=StorageType.Map, staleness=1,
                                  map_capacity=1 << 16)
        g = torch.Generator().manual_seed(0)
        keys = torch.sort(torch.unique(torch.randint(0, (1 << 32) - 1, (20_000,), generator=g)))[0]
        vals = torch.randn(keys.numel(), generator=g)
        out = {}

        def fn(info):
            t = info.create_kv_client_table(tid)
            t.add(keys, vals)
            t.add(keys[::2], vals[::2])
            t.clock()
            out["got"] = t.get(keys).cpu()
            out["missing"] = t.get(torch.tensor([5, 77, (1 << 31) + 3])).cpu()

        task = MLTask()
        task.set_worker_alloc([WorkerAlloc(0, 1)])
        task.set_tables([tid])
        task.set_lambda(fn)
        engine.run(task)
        ref = vals.clone()
        ref[::2] *= 2
        torch.testing.assert_close(out["got"], ref)
        assert out["missing"].abs().sum().item() == 0  # operator[] of an absent key reads 0
        table = engine.tables[tid]
        total = 0
        for s in table.own_shards:
            k, v = table.hash_items(s)
            lo, hi = table.ranges[s]
            assert bool(((k >= lo) & (k < hi)).all())
            total += k.numel()
        assert total == keys.numel()
    finally:
        engine.stop_everything()


@pytest.mark.parametrize("features", [50_000, 1_000_000, 5_000_000])
def test_packed_batch_loader_matches_plain_csr(features):
    """Compact H2D format (u8 row lengths, 16/20/24-bit ids, u8 labels) unpacks to exactly the CSR the plain
    loader stages; rows of varying length make the windows start at odd nnz offsets (bit offset 4 at 20 bits)."""
    from minips_b200.models.lr import CsrBatchSource
    from minips_b200.utils.loader import AsyncBatchLoader, PackedCsrHost
    g = torch.Generator().manual_seed(4)
    n = 20_000
    lens = torch.randint(1, 60, (n,), generator=g)
    row_ptr = torch.zeros(n + 1, dtype=torch.int64)
    row_ptr[1:] = torch.cumsum(lens, 0)
    nnz = int(row_ptr[-1])
    data = CsrBatchSource(row_ptr, torch.randint(0, features, (nnz,), generator=g, dtype=torch.int32),
                          torch.randn(nnz, generator=g), torch.randint(0, 2, (n,), generator=g).float())
    packed = PackedCsrHost.from_csr(data)
    assert packed is not None and packed.id_bits == {50_000: 16, 1_000_000: 20, 5_000_000: 24}[features]
    B = 4096
    plain = AsyncBatchLoader(data.pin(), B, "cuda")
    comp = AsyncBatchLoader(packed, B, "cuda")
    for start in (0, 1, 1234, 7777, n - B):
        a = plain.acquire(plain.prefetch(start))
        b = comp.acquire(comp.prefetch(start))
        torch.cuda.synchronize()
        m = int(row_ptr[start + B] - row_ptr[start])
        assert torch.equal(a.row_ptr - a.row_ptr[0], b.row_ptr)
        assert torch.equal(a.col[:m], b.col[:m])
        assert torch.equal(a.val[:m], b.val[:m])
        assert torch.equal(a.label, b.label)
    assert comp.bytes_per_batch < 0.87 * plain.bytes_per_batch
