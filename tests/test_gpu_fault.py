"""Fault tolerance on the device backend, proven by killing a process (needs 2 GPUs): master + two
ranks of minips_b200.apps.lr_example in fault-tolerant mode, checkpoint every 100 iterations, SIGKILL one
rank while the run is held at iteration 150; the master detects it after 3 silent heartbeat intervals,
relaunches it with --use_weight_file, orders the rollback; both ranks rebuild their engine in a new
generation, restore shards + clocks + iteration from the checkpoint of iteration 100 and run to 300.
The final weights equal those of an uninterrupted run that was itself restarted from its checkpoint of
iteration 100."""
import json
import os
import signal
import subprocess
import sys
import time

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rank_argv(hostfile, store_port, prefix, report, weights, extra):
    return [sys.executable, "-m", "minips_b200.apps.lr_example", f"--config_file={hostfile}",
            "--heartbeat_interval=1", f"--store_port={store_port}", "--kModelType=BSP", "--dtype=fp64",
            "--num_dims=20000", "--num_samples=20000", "--nnz_per_row=30", "--batch_size=2048", "--alpha=0.05",
            "--num_iters=300", "--checkpoint_toggle=true", "--checkpoint_interval=100",
            f"--checkpoint_file_prefix={prefix}", f"--report_prefix={report}", f"--weights_out={weights}",
            "--num_servers_per_node=2", "--seed=3"] + extra


def _run_cluster(tmp, tag, base, kill: bool, extra, preset_ckpt=None):
    d = tmp / tag
    d.mkdir()
    hostfile = d / "nodes"
    hostfile.write_text(f"0:localhost:{base}\n1:localhost:{base + 1}\n2:localhost:{base + 2}\n")
    prefix, report, weights = str(d / "ck_"), str(d / "rep"), str(d / "w")
    if preset_ckpt:
        for f in os.listdir(preset_ckpt):
            if f.startswith("ck_"):
                (d / f).write_bytes((preset_ckpt / f).read_bytes())
    launch = d / "launch.json"
    launch.write_text(json.dumps({
        "argv": _rank_argv(hostfile, base + 7, prefix, report, weights, ["--my_id={id}"] + extra),
        "log": str(d / "relaunch_{id}.log"), "device_of": {"0": 0, "2": 1}}))
    env = dict(os.environ, MINIPS_LOG="0", MINIPS_SPIN_TIMEOUT_S="5")
    master = subprocess.Popen(
        [sys.executable, "-m", "minips_b200.apps.master", f"--config_file={hostfile}", "--heartbeat_interval=1",
         f"--relaunch_cmd={sys.executable} {os.path.join(ROOT, 'scripts', 'relaunch_rank.py')} {launch}",
         f"--store_port={base + 7}"], cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    time.sleep(1.0)
    procs = {}
    for i, devi in ((0, 0), (2, 1)):
        procs[i] = subprocess.Popen(
            _rank_argv(hostfile, base + 7, prefix, report, weights, [f"--my_id={i}"] + extra), cwd=ROOT,
            env=dict(env, LOCAL_RANK=str(devi)), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    outs = {}
    try:
        if kill:
            end = time.time() + 180
            while time.time() < end and not os.path.exists(f"{report}_paused_2"):
                assert procs[2].poll() is None, procs[2].communicate()[0]
                time.sleep(0.1)
            assert os.path.exists(f"{report}_paused_2"), "rank 2 never reached the pause at iteration 150"
            procs[2].send_signal(signal.SIGKILL)
            procs[2].wait()
            outs[2] = ""
        for i, p in procs.items():
            if i not in outs:
                outs[i], _ = p.communicate(timeout=300)
                assert p.returncode == 0, outs[i][-4000:]
        mout, _ = master.communicate(timeout=120)
        assert master.returncode == 0, mout[-3000:]
    finally:
        for p in list(procs.values()) + [master]:
            if p.poll() is None:
                p.kill()
    return d, outs, mout


def test_kill_a_rank_resume_equals_uninterrupted(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    base = 24000 + os.getpid() % 2000
    # A: killed at iteration 150, recovered from the checkpoint of iteration 100
    da, outs, mout = _run_cluster(tmp_path, "killed", base, True, ["--pause_at_iter=150", "--pause_s=120"])
    relog = (da / "relaunch_2.log").read_text()
    assert "[Fault Tolerance][Phase2]" in mout and "[Fault Tolerance][Phase4]" in mout, mout
    assert "[Fault Tolerance][Phase3]" in relog and "[Fault Tolerance][Phase5]" in relog, relog[-3000:]
    assert "[Fault Tolerance][Phase5]" in outs[0], outs[0][-3000:]
    assert "resumed from checkpoint at iteration 100" in outs[0] and "resumed from checkpoint at iteration 100" in relog
    # B: uninterrupted to 100 (B1), then a fresh cluster restarted from that checkpoint to 300 (B2)
    db1, _, _ = _run_cluster(tmp_path, "first100", base + 20, False, ["--num_iters=100"])
    db2, _, _ = _run_cluster(tmp_path, "restart", base + 40, False, ["--use_weight_file=true"], preset_ckpt=db1)
    wa = torch.load(da / "w_0.pt")
    wb = torch.load(db2 / "w_0.pt")
    assert wa.dtype == torch.float64 and wa.abs().max() > 0
    torch.testing.assert_close(wa, wb, rtol=1e-9, atol=1e-12)
    torch.testing.assert_close(torch.load(da / "w_2.pt"), wa, rtol=0, atol=0)   # every rank gathers the same table
    # evidence for profiles/: the stamps of the run, in the order they were written
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        keep = ("[Fault Tolerance]", "checkpoint at iteration", "resumed from checkpoint", "generation",
                "rollbacks ordered", "Total training time", "supervising")
        with open(os.path.join(out_dir, "ft_gpu_kill_test.log"), "w") as f:
            for name, text in (("master", mout), ("rank 0 (survivor)", outs[0]), ("rank 2 (relaunched)", relog)):
                f.write(f"==== {name}\n")
                f.writelines(l + "\n" for l in text.splitlines() if any(k in l for k in keep))
            f.write(f"==== max |w_killed_and_recovered - w_uninterrupted| = {float((wa - wb).abs().max()):.3e} "
                    f"(max |w| = {float(wa.abs().max()):.3e})\n")
