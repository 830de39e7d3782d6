"""Elastic scale-out on the device backend (needs 2 GPUs; reference lr_example.cpp:396-409,
master/master_thread.cpp:40-43, comm/mailbox.cpp:197-219): a job of ONE rank (GPU 0) trains with a
checkpoint every 100 iterations and is held at iteration 150; a new process (GPU 1) joins with --scale.
The master adds it and orders kScaleRollback; both ranks meet in generation 1, the table is re-created over
BOTH GPUs (the restore re-shards the one-node checkpoint), both resume at iteration 100 and run to 300.
The final weights equal those of a plain two-rank run started from a copy of the same checkpoint."""
import os
import shutil
import subprocess
import sys
import time

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

COMMON = ["--kModelType=BSP", "--dtype=fp64", "--num_dims=20000", "--num_samples=20000", "--nnz_per_row=30",
          "--batch_size=2048", "--alpha=0.05", "--num_iters=300", "--checkpoint_toggle=true",
          "--checkpoint_interval=100", "--num_servers_per_node=2", "--seed=3"]


def test_scale_out_resharded_resume_equals_two_rank_run(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    base = 28000 + os.getpid() % 1400
    d = tmp_path / "elastic"
    d.mkdir()
    hostfile = d / "nodes"
    hostfile.write_text(f"0:localhost:{base}\n1:localhost:{base + 1}\n")
    prefix, report, weights, scale_file = str(d / "ck_"), str(d / "rep"), str(d / "w"), str(d / "scale")
    env = dict(os.environ, MINIPS_LOG="0", MINIPS_SPIN_TIMEOUT_S="5")
    ft = [f"--config_file={hostfile}", "--heartbeat_interval=1", f"--store_port={base + 7}",
          f"--checkpoint_file_prefix={prefix}", f"--report_prefix={report}", f"--weights_out={weights}",
          f"--scale_file={scale_file}"]
    app = [sys.executable, "-m", "minips_b200.apps.lr_example"]
    master = subprocess.Popen(
        [sys.executable, "-m", "minips_b200.apps.master", f"--config_file={hostfile}", "--heartbeat_interval=1",
         f"--store_port={base + 7}", f"--scale_file={scale_file}"], cwd=ROOT, env=env, stdout=subprocess.PIPE,
        stderr=subprocess.STDOUT, text=True)
    time.sleep(1.0)
    procs = {0: subprocess.Popen(app + COMMON + ft + ["--my_id=0", "--pause_at_iter=150", "--pause_s=150"], cwd=ROOT,
                                 env=dict(env, LOCAL_RANK="0"), stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                 text=True)}
    snap = tmp_path / "snapshot"
    snap.mkdir()
    outs = {}
    try:
        end = time.time() + 180
        while time.time() < end and not os.path.exists(f"{report}_paused_0"):
            assert procs[0].poll() is None, procs[0].communicate()[0]
            time.sleep(0.1)
        assert os.path.exists(f"{report}_paused_0"), "rank 0 never reached the pause at iteration 150"
        for f in os.listdir(d):                       # the checkpoint of iteration 100, before anyone overwrites it
            if f.startswith("ck_"):
                shutil.copy(d / f, snap / f)
        procs[2] = subprocess.Popen(
            # the same flags as rank 0 (the pause only acts in a first generation): every rank must cut the run
            # into the same launches, the exchange kernel counts its hand-overs per launch
            app + COMMON + ft + ["--my_id=2", "--scale=true", "--scale_name=localhost", f"--scale_port={base + 2}",
                                 "--pause_at_iter=150", "--pause_s=150"],
            cwd=ROOT, env=dict(env, LOCAL_RANK="1"), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        for i, p in procs.items():
            outs[i], _ = p.communicate(timeout=300)
            assert p.returncode == 0, outs[i][-4000:]
        mout, _ = master.communicate(timeout=120)
        assert master.returncode == 0, mout[-3000:]
    finally:
        for p in list(procs.values()) + [master]:
            if p.poll() is None:
                p.kill()
    assert "scale-outs for [2]" in mout, mout
    for i in (0, 2):
        assert "resumed from checkpoint at iteration 100" in outs[i], outs[i][-3000:]
    assert "ck_server_layout" in os.listdir(snap)
    # baseline: two ranks under torchrun, started from the snapshot of the ONE-node checkpoint
    b = tmp_path / "baseline"
    b.mkdir()
    for f in os.listdir(snap):
        shutil.copy(snap / f, b / f)
    p = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr=127.0.0.1",
         f"--master-port={base + 30}", "-m", "minips_b200.apps.lr_example"] + COMMON +
        ["--use_weight_file=true", f"--checkpoint_file_prefix={b / 'ck_'}", f"--weights_out={b / 'w'}"],
        cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    wa, wb = torch.load(d / "w_0.pt"), torch.load(b / "w_0.pt")
    assert wa.dtype == torch.float64 and wa.abs().max() > 0
    torch.testing.assert_close(wa, wb, rtol=1e-9, atol=1e-12)
    torch.testing.assert_close(torch.load(d / "w_2.pt"), wa, rtol=0, atol=0)
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        keep = ("scaling out", "checkpoint at iteration", "resumed from checkpoint", "generation", "scale-outs",
                "Total training time", "supervising", "[Fault Tolerance]")
        with open(os.path.join(out_dir, "elastic_scale_out_test.log"), "w") as f:
            for name, text in (("master", mout), ("rank 0 (was alone)", outs[0]), ("node 2 (scaled in)", outs[2])):
                f.write(f"==== {name}\n")
                f.writelines(l + "\n" for l in text.splitlines() if any(k in l for k in keep))
            f.write(f"==== max |w_scaled_out - w_two_rank_run_from_the_same_checkpoint| = "
                    f"{float((wa - wb).abs().max()):.3e} (max |w| = {float(wa.abs().max()):.3e})\n")
