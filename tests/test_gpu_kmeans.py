"""tcgen05 k-means assignment + the mini-batch update against fp32 PyTorch oracles.

The tensor-core path rounds nothing beyond the bf16 inputs but accumulates in a different
order than fp32 matmul and packs the arg-min index into 7 low mantissa bits, so near-ties may
legitimately pick a different centre: assignments are judged by OBJECTIVE (the chosen centre's
true distance is within eps of the true minimum), not by index equality.
"""
import pytest
import torch

from minips_b200.parallel.types import MLTask, ModelType, Node, WorkerAlloc

pytestmark = pytest.mark.gpu


def _check_objective(assign, dist, d_full, frac_exact=0.98):
    dmin = d_full.min(1).values
    scale = d_full.abs().max().item()
    tol = 2e-3 * scale + 1e-3
    chosen = d_full.gather(1, assign.long()[:, None])[:, 0]
    assert assign.min().item() >= 0 and assign.max().item() < d_full.shape[1]
    assert (chosen - dmin).max().item() <= tol, ((chosen - dmin).max().item(), tol)
    assert (dist - dmin).abs().max().item() <= tol, ((dist - dmin).abs().max().item(), tol)
    exact = (assign.long() == d_full.argmin(1)).float().mean().item()
    assert exact >= frac_exact, exact


@pytest.mark.parametrize("n,K,D", [(5000, 256, 128), (384, 128, 128), (100_003, 1024, 128), (7777, 384, 64)])
def test_assign_tc_matches_oracle(n, K, D):
    from minips_b200.models.kmeans import KMeansAssigner, oracle_assign
    g = torch.Generator(device="cuda").manual_seed(n + K)
    x = torch.randn(n, D, device="cuda", generator=g).to(torch.bfloat16)
    centers = torch.randn(K, D, device="cuda", generator=g)
    km = KMeansAssigner(K, D, "cuda")
    assign, dist = km.assign(x, centers=centers)
    torch.cuda.synchronize()
    _, _, d_full = oracle_assign(x, centers)
    _check_objective(assign, dist, d_full)
    # the fused prologue left the rounded centres, their norms and the fp32 copy behind
    torch.testing.assert_close(km.c_bf16[:K * D].view(K, D).float(), -2.0 * centers.to(torch.bfloat16).float())
    cf = centers.to(torch.bfloat16).float()
    torch.testing.assert_close(km.cnorm, (cf * cf).sum(1), rtol=1e-5, atol=1e-4)
    # ||c||^2 rides into the accumulator as hi + mid + lo (bf16) through one extra MMA step
    aug = km.c_bf16[K * D:K * D + K * 64].view(K, 64).float()
    torch.testing.assert_close(aug[:, :3].sum(1), km.cnorm, rtol=1e-6, atol=1e-5)
    assert float(aug[:, 3:].abs().sum()) == 0.0
    torch.testing.assert_close(km.c_f32.view(K, D), centers)
    # differential: plain CUDA-core kernel on the same rounded inputs
    a_ref, d_ref = km.assign_ref(x[:2048])
    _check_objective(a_ref, d_ref, d_full[:2048], frac_exact=0.999)


def test_assign_tc_few_ctas_many_groups():
    """Persistent loop: 2 CTAs walk 40 groups (ring phases wrap many times)."""
    from minips_b200.models.kmeans import KMeansAssigner, oracle_assign
    n, K, D = 40 * 384 + 17, 256, 128
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(n, D, device="cuda", generator=g).to(torch.bfloat16)
    centers = torch.randn(K, D, device="cuda", generator=g)
    km = KMeansAssigner(K, D, "cuda")
    assign, dist = km.assign(x, centers=centers, max_ctas=2)
    _, _, d_full = oracle_assign(x, centers)
    _check_objective(assign, dist, d_full)


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("model", [ModelType.BSP, ModelType.SSP])
def test_kmeans_training_matches_oracle(model, fused):
    from minips_b200.models.kmeans import KMeansTrainer, init_random, make_blobs, oracle_step
    from minips_b200.parallel.engine import Engine
    K, D, n, batch, iters, alpha = 128, 128, 40_000, 8192, 4, 1.0
    node = Node(0, "localhost", 0)
    engine = Engine(node, [node], device=0)
    engine.start_everything(num_server_threads_per_node=8)
    try:
        x = make_blobs(n, D, K, seed=3, spread=0.05)
        t_c = engine.create_table(num_dims=K * D, model_type=model, staleness=1)
        t_n = engine.create_table(num_dims=K, model_type=model, staleness=1)
        c0 = init_random(x, K, seed=1)
        out = {}

        def fn(info):
            tc, tn = info.create_kv_client_table(t_c), info.create_kv_client_table(t_n)
            # "set" = add onto zero (kmeans.cpp:185), then clock so the values are visible under BSP
            tc.add_all(c0.reshape(-1).contiguous())
            tc.clock()
            tn.clock()
            tr = KMeansTrainer(tc, tn, x, K, alpha)
            tr.allow_fused = fused
            assert tr._fusable() == fused
            sses = [tr.step(i * batch, batch).clone() for i in range(iters)]
            out["centers"] = tc.get_all().view(K, D)
            out["counts"] = tn.get_all()
            out["sse"] = torch.cat(sses).cpu()

        task = MLTask()
        task.set_worker_alloc([WorkerAlloc(0, 1)])
        task.set_tables([t_c, t_n])
        task.set_lambda(fn)
        engine.run(task)
    finally:
        engine.stop_everything()
    c, cnt = c0.clone(), torch.zeros(K, device="cuda")
    ref_sse = []
    for i in range(iters):
        c, cnt, s = oracle_step(c, cnt, [x[i * batch:(i + 1) * batch]], alpha)
        ref_sse.append(s)
    from minips_b200.models.kmeans import _min_sqdist_torch
    assert out["counts"].sum().item() == iters * batch
    # Random seeding puts several centres into the same blob, so exact ties are common and the
    # two implementations may split a blob's points differently: judge by objective.  The first
    # iteration (identical centres on both sides) must agree tightly; later ones within 1 %.
    assert abs(out["sse"][0].item() - ref_sse[0]) <= 2e-3 * ref_sse[0] + 1.0
    for a, b in zip(out["sse"].tolist(), ref_sse):
        assert abs(a - b) <= 1e-2 * abs(b) + 1.0, (a, b)
    assert out["sse"][-1] < out["sse"][0]
    cost_dev = _min_sqdist_torch(x, out["centers"]).mean().item()
    cost_ref = _min_sqdist_torch(x, c).mean().item()
    assert abs(cost_dev - cost_ref) <= 2e-2 * cost_ref + 1e-3, (cost_dev, cost_ref)
    assert (out["counts"] - cnt).abs().sum().item() <= 0.05 * iters * batch


def test_init_modes_cover_blobs():
    from minips_b200.models.kmeans import INIT_MODES, _min_sqdist_torch, make_blobs
    x = make_blobs(20_000, 64, 16, seed=9, spread=0.05)
    base = None
    for name, fn in INIT_MODES.items():
        c = fn(x, 16, seed=2)
        assert c.shape == (16, 64)
        cost = _min_sqdist_torch(x, c).mean().item()
        if name == "random":
            base = cost
        else:
            assert cost <= base * 1.05, (name, cost, base)  # D^2 seeding is no worse than random


@pytest.mark.parametrize("D", [128, 64])
def test_sorted_accumulate_matches_oracle(D):
    """Counting-sort + run-sum accumulation against plain torch index_add (fp32) and the atomic kernel."""
    import ctypes as C
    from minips_b200.ops import _lib
    lib = _lib.load()
    K, n = 256, 100_003
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(n, D, device="cuda", generator=g).to(torch.bfloat16)
    assign = torch.randint(0, K, (n,), device="cuda", generator=g, dtype=torch.int32)
    assign[assign == 7] = 8            # an empty centre
    assign[assign % 50 == 3] = 3       # a heavy one
    dist = torch.randn(n, device="cuda", generator=g)
    xf = x.float()
    ref_sums = torch.zeros(K, D, device="cuda").index_add_(0, assign.long(), xf)
    ref_cnt = torch.bincount(assign.long(), minlength=K).float()
    ref_sse = ((xf * xf).sum(1) + dist).clamp_min(0).sum()
    sp = _lib.stream_ptr()
    scratch = torch.zeros(3 * K + 1 + n, dtype=torch.int32, device="cuda")
    for rounds in (1, 2):              # twice: the scratch must come back reusable
        sums = torch.zeros(K * D, device="cuda")
        cnt = torch.zeros(K, device="cuda")
        sse = torch.zeros(1, device="cuda")
        _lib.check(lib.mps_kmeans_accumulate_sorted(
            C.c_void_p(x.data_ptr()), C.c_void_p(assign.data_ptr()), C.c_void_p(dist.data_ptr()), n, K, D,
            C.c_void_p(sums.data_ptr()), C.c_void_p(cnt.data_ptr()), C.c_void_p(sse.data_ptr()),
            C.c_void_p(scratch.data_ptr()), sp), "accumulate_sorted")
        torch.cuda.synchronize()
        assert torch.equal(cnt, ref_cnt)
        assert torch.allclose(sums.view(K, D), ref_sums, rtol=1e-4, atol=1e-2)
        assert abs(sse.item() - ref_sse.item()) <= 1e-4 * abs(ref_sse.item())
        assert int(scratch[:K].abs().sum()) == 0
    perm = scratch[3 * K + 1:3 * K + 1 + n].long()
    assert torch.equal(torch.sort(perm).values, torch.arange(n, device="cuda"))
    assert bool((assign[perm][1:] >= assign[perm][:-1]).all())


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("K,D", [(5, 123), (37, 20), (130, 64)])
def test_kmeans_reference_shapes(K, D, fused):
    """Any K and any num_dims <= 128 (the reference's own defaults are K=5, num_dims=123,
    scripts/kmeans.py:17-43): the tables hold exactly K*D and K keys, the tcgen05 kernel runs the
    problem padded (dummy centres with norm 2^127, zero columns)."""
    from minips_b200.models.kmeans import (KMeansTrainer, _min_sqdist_torch, init_random, make_blobs, oracle_step,
                                           pad_points, padded_shape)
    from minips_b200.parallel.engine import Engine
    n, batch, iters, alpha = 12_288, 4096, 3, 1.0
    node = Node(0, "localhost", 0)
    engine = Engine(node, [node], device=0)
    engine.start_everything(num_server_threads_per_node=3)
    try:
        x = make_blobs(n, D, K, seed=3, spread=0.05)              # [n, D] bf16, unpadded
        xp = pad_points(x)
        assert xp.shape[1] == padded_shape(K, D)[1]
        t_c = engine.create_table(num_dims=K * D, model_type=ModelType.BSP)
        t_n = engine.create_table(num_dims=K, model_type=ModelType.BSP)
        c0 = init_random(x, K, seed=1)
        out = {}

        def fn(info):
            tc, tn = info.create_kv_client_table(t_c), info.create_kv_client_table(t_n)
            tc.add_all(c0.reshape(-1).contiguous())
            tc.clock()
            tn.clock()
            tr = KMeansTrainer(tc, tn, xp, K, alpha, D=D)
            tr.allow_fused = fused
            sses = [tr.step(i * batch, batch).clone() for i in range(iters)]
            out["centers"] = tc.get_all().view(K, D)
            out["counts"] = tn.get_all()
            out["sse"] = torch.cat(sses).cpu()
            out["assign_max"] = int(tr._out[0].max())

        task = MLTask()
        task.set_worker_alloc([WorkerAlloc(0, 1)])
        task.set_tables([t_c, t_n])
        task.set_lambda(fn)
        engine.run(task)
    finally:
        engine.stop_everything()
    assert out["assign_max"] < K, "a padding centre was chosen"
    c, cnt = c0.clone(), torch.zeros(K, device="cuda")
    ref_sse = []
    for i in range(iters):
        c, cnt, s = oracle_step(c, cnt, [x[i * batch:(i + 1) * batch]], alpha)
        ref_sse.append(s)
    assert out["counts"].sum().item() == iters * batch
    assert abs(out["sse"][0].item() - ref_sse[0]) <= 2e-3 * ref_sse[0] + 1.0
    cost_dev = _min_sqdist_torch(x, out["centers"]).mean().item()
    cost_ref = _min_sqdist_torch(x, c).mean().item()
    assert abs(cost_dev - cost_ref) <= 2e-2 * cost_ref + 1e-3, (cost_dev, cost_ref)
    assert (out["counts"] - cnt).abs().sum().item() <= 0.05 * iters * batch


def test_kmeans_fused_with_two_workers_per_gpu():
    """Two workers share the GPU (the reference's default layout is 2 workers per node): each keeps the
    one-call fused step; their BSP dependencies are stream events, the last one to clock folds."""
    from minips_b200.models.kmeans import KMeansTrainer, _min_sqdist_torch, init_random, make_blobs, oracle_step
    from minips_b200.parallel.engine import Engine
    K, D, n, batch, iters, alpha = 128, 64, 32_768, 4096, 3, 1.0
    node = Node(0, "localhost", 0)
    engine = Engine(node, [node], device=0)
    engine.start_everything(num_server_threads_per_node=4)
    try:
        x = make_blobs(n, D, K, seed=5, spread=0.05)
        halves = [x[: n // 2], x[n // 2:]]
        t_c = engine.create_table(num_dims=K * D, model_type=ModelType.BSP)
        t_n = engine.create_table(num_dims=K, model_type=ModelType.BSP)
        c0 = init_random(x, K, seed=1)
        out = {}

        def fn(info):
            tc, tn = info.create_kv_client_table(t_c), info.create_kv_client_table(t_n)
            if info.worker_id == 0:
                tc.add_all(c0.reshape(-1).contiguous())
            tc.clock()
            tn.clock()
            tr = KMeansTrainer(tc, tn, halves[info.worker_id], K, alpha)
            assert tr._fusable()
            for i in range(iters):
                tr.step(i * batch, batch)
            if info.worker_id == 0:
                out["centers"] = tc.get_all().view(K, D).clone()
                out["counts"] = tn.get_all().clone()

        task = MLTask()
        task.set_worker_alloc([WorkerAlloc(0, 2)])
        task.set_tables([t_c, t_n])
        task.set_lambda(fn)
        engine.run(task)
    finally:
        engine.stop_everything()
    c, cnt = c0.clone(), torch.zeros(K, device="cuda")
    for i in range(iters):
        c, cnt, _ = oracle_step(c, cnt, [h[i * batch:(i + 1) * batch] for h in halves], alpha)
    assert out["counts"].sum().item() == 2 * iters * batch
    cost_dev = _min_sqdist_torch(x, out["centers"]).mean().item()
    cost_ref = _min_sqdist_torch(x, c).mean().item()
    assert abs(cost_dev - cost_ref) <= 2e-2 * cost_ref + 1e-3, (cost_dev, cost_ref)
    assert (out["counts"] - cnt).abs().sum().item() <= 0.05 * 2 * iters * batch


def test_kmeans_app_on_libsvm_input(tmp_path):
    """scripts/kmeans.py-style run: libsvm file in, K=5, num_dims=123 (the reference's local defaults)."""
    import subprocess
    import sys
    import os
    g = torch.Generator().manual_seed(2)
    rows = []
    centers = torch.rand(5, 123, generator=g)
    for i in range(3000):
        c = centers[i % 5]
        on = torch.nonzero(torch.rand(123, generator=g) < c * 0.3)[:, 0]
        rows.append(f"{i % 2} " + " ".join(f"{int(j) + 1}:1" for j in on.tolist()))
    path = tmp_path / "a9a_like.txt"
    path.write_text("\n".join(rows) + "\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-m", "minips_b200.apps.kmeans", f"--input={path}", "--K=5", "--num_dims=123",
                        "--batch_size=512", "--num_iters=12", "--kModelType=SSP", "--kStaleness=1",
                        "--kmeans_init_mode=kmeans++", "--report_interval=4", "--alpha=0.5"],
                       cwd=root, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-3000:]
    sse = [float(l.split("SSE/point")[1].split()[0]) for l in p.stdout.splitlines() if "SSE/point" in l]
    assert len(sse) == 3 and sse[-1] <= sse[0] * 1.05, p.stdout
    assert "loaded 3000 points" in p.stdout
