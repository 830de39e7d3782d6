"""Dense LR: tcgen05 forward (+ fused pull / sigmoid epilogue) and the gradient / push path
against an fp32 PyTorch oracle on the same bf16-rounded inputs."""
import pytest
import torch

from minips_b200.parallel.types import MLTask, ModelType, Node, WorkerAlloc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["rows", "stream", "mma", "twopass"])
@pytest.mark.parametrize("n,D", [(4096, 512), (10_007, 1000), (300, 64), (2048, 4096), (3001, 1024), (29, 8)])
def test_dense_lr_training_matches_oracle(n, D, mode, monkeypatch):
    """Four implementations of the iteration against the same oracle: "rows" (warp per row, bulk-copy ring,
    nothing resident; the default), "stream" (TMA-fed, 64-row tile resident), "mma" (both products on tcgen05,
    the second one with the tile as an MN-major operand) and "twopass" (tcgen05 forward + CUDA-core gradient
    pass; also what D > 1024 takes)."""
    monkeypatch.setenv("MINIPS_DENSE_LR", mode)
    from minips_b200.models.dense_lr import DenseLRTrainer, oracle_step
    from minips_b200.parallel.engine import Engine
    node = Node(0, "localhost", 0)
    engine = Engine(node, [node], device=0)
    engine.start_everything(num_server_threads_per_node=4)
    try:
        g = torch.Generator(device="cuda").manual_seed(n + D)
        x = (torch.randn(n, D, device="cuda", generator=g) / D ** 0.5).to(torch.bfloat16)
        w_true = torch.randn(D, device="cuda", generator=g) * 4
        y = (torch.rand(n, device="cuda", generator=g) < torch.sigmoid(x.float() @ w_true)).float()
        tid = engine.create_table(num_dims=D, model_type=ModelType.BSP)
        alpha, iters = 0.5, 4
        out = {}

        def fn(info):
            tr = DenseLRTrainer(info.create_kv_client_table(tid), x, y, alpha)
            losses = [tr.step(0, n).clone() for _ in range(iters)]
            out["w"] = tr.client.get_all()
            out["loss"] = torch.stack(losses).cpu()

        task = MLTask()
        task.set_worker_alloc([WorkerAlloc(0, 1)])
        task.set_tables([tid])
        task.set_lambda(fn)
        engine.run(task)
    finally:
        engine.stop_everything()
    w = torch.zeros(D, device="cuda")
    for i in range(iters):
        d, loss, correct = oracle_step(w, x, y, alpha)
        assert abs(out["loss"][i, 0].item() - loss) <= 2e-3 * abs(loss) + 1e-2, (i, out["loss"][i, 0].item(), loss)
        assert abs(out["loss"][i, 1].item() - correct) <= max(2, 0.002 * n)
        w = w + d
    torch.testing.assert_close(out["w"], w, rtol=5e-3, atol=2e-3)
    assert out["loss"][-1, 0] < out["loss"][0, 0]
