"""The C++ applications on the host (CPU) runtime, launched the way scripts/*.py launches them:
one OS process per hostfile line on localhost (BASELINE.json config 1: LR on a synthetic
10k x 1k libsvm file, localnodes-shaped hostfile, BSP, CPU only)."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "build", "bin")


@pytest.fixture(scope="module")
def built():
    sys.path.insert(0, ROOT)
    import importlib.util
    spec = importlib.util.spec_from_file_location("minips_build", os.path.join(ROOT, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build_core(verbose=False)
    return BIN


@pytest.fixture(scope="module")
def dataset(tmp_path_factory):
    from minips_b200.ops.sparse import write_synthetic_libsvm
    d = tmp_path_factory.mktemp("data")
    return write_synthetic_libsvm(str(d / "cfg1.libsvm"), 10_000, 1000, 10, seed=3)


def _hostfile(tmp_path, n, base):
    p = tmp_path / "nodes"
    p.write_text("".join(f"{i}:localhost:{base + i}\n" for i in range(n)))
    return str(p)


def _launch(exe, hostfile, n, extra, timeout=180):
    procs = [subprocess.Popen([exe, f"--my_id={i}", f"--config_file={hostfile}"] + extra, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for i in range(n)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    return [p.returncode for p in procs], outs


def test_basic_example_three_nodes(built, tmp_path):
    base = 21000 + os.getpid() % 2000
    hf = _hostfile(tmp_path, 3, base)
    rcs, outs = _launch(os.path.join(built, "basic_example"), hf, 3, ["--num_iters=30", "--num_workers_per_node=4"])
    assert rcs == [0, 0, 0], outs
    assert "every key == 180: true" in outs[0]  # 0.5 * 30 iterations * 4 workers * 3 nodes


@pytest.mark.parametrize("model", ["BSP", "SSP", "ASP"])
def test_lr_config1_localnodes(built, dataset, tmp_path, model):
    base = 23000 + os.getpid() % 2000 + {"BSP": 0, "SSP": 10, "ASP": 20}[model]
    hf = _hostfile(tmp_path, 3, base)
    ck = str(tmp_path / "ck_")
    extra = [f"--input={dataset}", "--num_dims=1000", f"--kModelType={model}", "--kStaleness=2", "--batch_size=100",
             "--num_iters=150", "--num_workers_per_node=2", "--alpha=0.05", "--heartbeat_interval=-1",
             "--block_size=8192", f"--assigner_master_port={base + 5}", "--checkpoint_toggle=true",
             f"--checkpoint_file_prefix={ck}", "--checkpoint_interval=100"]
    rcs, outs = _launch(os.path.join(built, "lr_example"), hf, 3, extra)
    assert rcs == [0, 0, 0], outs
    loaded = [int(re.search(r"loaded (\d+) samples", o).group(1)) for o in outs]
    assert sum(loaded) == 10_000 and min(loaded) > 0  # dynamic block assignment covers the file exactly once
    accs = [float(m) for o in outs for m in re.findall(r"accuracy ([0-9.]+)", o)]
    assert len(accs) == 3 and min(accs) > 0.85, accs
    # the checkpoint at iteration 100 returned (it hangs under BSP / ASP in the reference) and wrote the three files
    assert any("checkpoint at iteration 100" in o for o in outs)
    for i in range(3):
        assert os.path.exists(f"{ck}server_params_{i}") and os.path.exists(f"{ck}server_progress_{i}")
        assert os.path.exists(f"{ck}worker_config_{i}")
    line = open(f"{ck}server_params_0").read()
    assert "\n" not in line and re.match(r"^(\d+:-?[0-9.e+-]+ )+$", line)


def test_lr_with_master_heartbeat(built, dataset, tmp_path):
    """heartbeat_interval > 0: node 1 becomes the master (no servers / workers), nodes 0 and 2 train."""
    base = 25000 + os.getpid() % 2000
    hf = _hostfile(tmp_path, 3, base)
    extra = [f"--input={dataset}", "--num_dims=1000", "--kModelType=SSP", "--batch_size=50", "--num_iters=60",
             "--num_workers_per_node=2", "--heartbeat_interval=1", "--block_size=8192",
             f"--assigner_master_port={base + 5}"]
    rcs, outs = _launch(os.path.join(built, "lr_example"), hf, 3, extra)
    assert rcs == [0, 0, 0], outs
    assert "Total training time" in outs[0] and "Total training time" in outs[2]
    assert "Total training time" not in outs[1]  # the master only supervises


def test_kmeans_two_nodes(built, dataset, tmp_path):
    base = 27000 + os.getpid() % 2000
    hf = _hostfile(tmp_path, 2, base)
    extra = [f"--input={dataset}", "--num_dims=1000", "--K=4", "--kmeans_init_mode=kmeans_parallel", "--num_iters=6",
             "--batch_size=200", "--kModelType=BSP", "--block_size=8192", f"--assigner_master_port={base + 5}"]
    rcs, outs = _launch(os.path.join(built, "kmeans"), hf, 2, extra)
    assert rcs == [0, 0], outs
    assert "seeding took" in outs[0] and "SSE" in outs[0]
