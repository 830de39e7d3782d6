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
    ck = str(tmp_path / "ck_")
    common = ["--num_dims=50000", "--num_samples=40000", "--batch_size=4096", "--kModelType=BSP", "--alpha=0.05",
              "--num_servers_per_node=4", "--checkpoint_toggle=true", f"--checkpoint_file_prefix={ck}",
              "--checkpoint_interval=100"]
    out = _run("minips_b200.apps.lr_example", common + ["--num_iters=200"])
    assert "checkpoint at iteration 100" in out and "checkpoint at iteration 200" in out
    assert os.path.exists(ck + "server_params_0") and os.path.exists(ck + "server_params_0.3")
    assert os.path.exists(ck + "server_progress_0") and os.path.exists(ck + "worker_config_0")
    acc1 = float(out.split("accuracy ")[1].split()[0])
    # resume from the dump and keep training: starts at iteration 200
    out2 = _run("minips_b200.apps.lr_example", common + ["--num_iters=300", "--use_weight_file=true"])
    assert "resumed from checkpoint at iteration 200" in out2
    acc2 = float(out2.split("accuracy ")[1].split()[0])
    assert acc1 > 0.6 and acc2 >= acc1 - 0.02


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
