"""The Python applications (device backend) end to end on one GPU, as a user would launch them."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(module, args, timeout=240):
    env = dict(os.environ, MINIPS_SPIN_TIMEOUT_S="5", PYTHONPATH=ROOT)
    p = subprocess.run([sys.executable, "-m", module] + args, capture_output=True, text=True, timeout=timeout, cwd=ROOT,
                       env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-3000:]
    return p.stdout


def test_basic_example():
    out = _run("minips_b200.apps.basic_example", ["--workers=4", "--iters=20"])
    assert "every key == 40.0: True" in out


def test_lr_example_checkpoint_and_resume(tmp_path):
    """Resume equals the uninterrupted run: 300 iterations in one go vs 200, stop, resume from
    the checkpoint of iteration 200 (shards, clocks and iteration restored) and run to 300.  fp64 tables, so
    the two differ only by the order of floating-point atomics."""
    import shutil
    import torch
    ck_a, ck_b = str(tmp_path / "a_"), str(tmp_path / "b_")
    common = ["--num_dims=50000", "--num_samples=40000", "--batch_size=4096", "--kModelType=BSP", "--alpha=0.05",
              "--num_servers_per_node=4", "--checkpoint_toggle=true", "--checkpoint_interval=100", "--dtype=fp64"]
    out = _run("minips_b200.apps.lr_example", common + ["--num_iters=300", f"--checkpoint_file_prefix={ck_a}",
                                                       f"--weights_out={tmp_path / 'wa'}"])
    assert "checkpoint at iteration 100" in out and "checkpoint at iteration 300" in out
    out = _run("minips_b200.apps.lr_example", common + ["--num_iters=200", f"--checkpoint_file_prefix={ck_b}"])
    assert "checkpoint at iteration 100" in out and "checkpoint at iteration 200" in out
    assert os.path.exists(ck_b + "server_params_0") and os.path.exists(ck_b + "server_params_0.3")
    assert os.path.exists(ck_b + "server_progress_0") and os.path.exists(ck_b + "worker_config_0")
    assert "min_clock:200" in open(ck_b + "server_progress_0").read()
    acc1 = float(out.split("accuracy ")[1].split()[0])
    out2 = _run("minips_b200.apps.lr_example", common + ["--num_iters=300", "--use_weight_file=true",
                                                        f"--checkpoint_file_prefix={ck_b}",
                                                        f"--weights_out={tmp_path / 'wb'}"])
    assert "resumed from checkpoint at iteration 200" in out2
    acc2 = float(out2.split("accuracy ")[1].split()[0])
    assert acc1 > 0.6 and acc2 >= acc1 - 0.02
    wa, wb = torch.load(tmp_path / "wa_0.pt"), torch.load(tmp_path / "wb_0.pt")
    assert wa.dtype == torch.float64 and float(wa.abs().max()) > 0
    torch.testing.assert_close(wb, wa, rtol=1e-9, atol=1e-12)
    # the resumed run checkpointed again at 300: its tracker says so
    assert "min_clock:300" in open(ck_b + "server_progress_0").read()
    shutil.rmtree(tmp_path, ignore_errors=True)


def test_lr_example_checkpoint_with_two_workers_per_node(tmp_path):
    """Several workers in one process dump ONCE per process (they meet on a thread barrier first) and the
    resumed run starts where the dump says."""
    ck = str(tmp_path / "ck_")
    common = ["--num_dims=3000000", "--num_samples=20000", "--batch_size=512", "--kModelType=SSP", "--kStaleness=1",
              "--num_workers_per_node=2", "--num_servers_per_node=2", "--checkpoint_toggle=true",
              f"--checkpoint_file_prefix={ck}", "--checkpoint_interval=20"]
    out = _run("minips_b200.apps.lr_example", common + ["--num_iters=40"])
    assert "checkpoint at iteration 20" in out and "checkpoint at iteration 40" in out
    cfg = open(ck + "worker_config_0").read().split()
    assert sorted(cfg) == ["0:40", "1:40"], cfg
    assert "min_clock:40" in open(ck + "server_progress_0").read()
    out2 = _run("minips_b200.apps.lr_example", common + ["--num_iters=60", "--use_weight_file=true"])
    assert "resumed from checkpoint at iteration 40" in out2 and "checkpoint at iteration 60" in out2


def test_lr_example_sparse_regime_ssp():
    out = _run("minips_b200.apps.lr_example", ["--num_dims=5000000", "--num_samples=20000", "--batch_size=512",
                                                "--num_iters=50", "--kModelType=SSP", "--kStaleness=2"])
    assert "Total training time" in out


def test_lr_example_from_libsvm_file(tmp_path):
    from minips_b200.ops.sparse import write_synthetic_libsvm
    path = write_synthetic_libsvm(str(tmp_path / "a.libsvm"), 10_000, 1000, 10, seed=3)
    out = _run("minips_b200.apps.lr_example", [f"--input={path}", "--num_dims=1000", "--batch_size=500",
                                                "--num_iters=300", "--kModelType=BSP", "--alpha=0.02"])
    assert float(out.split("accuracy ")[1].split()[0]) > 0.85


def test_kmeans_app():
    out = _run("minips_b200.apps.kmeans", ["--K=128", "--num_dims=128", "--num_samples=200000", "--batch_size=32768",
                                           "--num_iters=10", "--kModelType=BSP", "--kmeans_init_mode=kmeans++",
                                           "--report_interval=5", "--alpha=1.0"])
    lines = [l for l in out.splitlines() if l.startswith("iter ")]
    assert len(lines) == 2
    first, last = (float(l.split("SSE/point ")[1].split()[0]) for l in lines)
    assert last <= first * 1.05
