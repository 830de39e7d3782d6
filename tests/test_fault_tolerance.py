"""Kill a real process and check the resumed run: master + two worker nodes as separate
processes, SIGKILL one worker in mid-run, the master (C++ Master / HeartBeatCheckThread through
minips_b200.core) detects it after 3 silent intervals, runs the relaunch command, orders the rollback
when the relaunched process reports back; both nodes resume from the last checkpoint in a new
generation and finish with exactly the result of an uninterrupted run.  CPU only: the data plane is a
toy (tests/ft_worker.py); the device version is tests/test_gpu_fault.py."""
import json
import os
import signal
import subprocess
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _wait_for(path, pred, timeout):
    end = time.time() + timeout
    while time.time() < end:
        try:
            if pred(open(path).read()):
                return True
        except (OSError, ValueError):
            pass
        time.sleep(0.05)
    return False


def test_kill_detect_relaunch_rollback_resume(tmp_path):
    base = 26000 + os.getpid() % 2000
    hostfile = tmp_path / "nodes"
    hostfile.write_text(f"0:localhost:{base}\n1:localhost:{base + 1}\n2:localhost:{base + 2}\n")
    out = tmp_path / "out"
    out.mkdir()
    args = {"config_file": str(hostfile), "heartbeat_interval": 1, "store_port": base + 7, "out_dir": str(out),
            "num_iters": 120, "checkpoint_interval": 20, "iter_sleep": 0.03}
    worker = os.path.join(ROOT, "tests", "ft_worker.py")
    launch = tmp_path / "launch.json"

    def argv_for(i):
        return [sys.executable, worker, json.dumps(dict(args, my_id=i))]

    launch.write_text(json.dumps({"argv": [sys.executable, worker, json.dumps(dict(args, my_id="{id}"))],
                                  "log": str(tmp_path / "relaunch_{id}.log")}))
    env = dict(os.environ, MINIPS_LOG="0")
    master = subprocess.Popen(
        [sys.executable, "-m", "minips_b200.apps.master", f"--config_file={hostfile}", "--heartbeat_interval=1",
         f"--relaunch_cmd={sys.executable} {os.path.join(ROOT, 'scripts', 'relaunch_rank.py')} {launch}",
         f"--store_port={base + 7}"], cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    time.sleep(1.0)
    procs = {i: subprocess.Popen(argv_for(i), cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                 text=True) for i in (0, 2)}
    try:
        assert _wait_for(out / "progress_2", lambda s: int(s) >= 50, 60), "node 2 never reached iteration 50"
        procs[2].send_signal(signal.SIGKILL)
        procs[2].wait()
        out0, _ = procs[0].communicate(timeout=120)
        assert procs[0].returncode == 0, out0
        mout, _ = master.communicate(timeout=60)
        assert master.returncode == 0, mout
    finally:
        for p in list(procs.values()) + [master]:
            if p.poll() is None:
                p.kill()
    relog = (tmp_path / "relaunch_2.log").read_text()
    # the five stamps of the reference's fault-tolerance timeline (report.md:90-94), phase 1 being the checkpoints
    assert "[Fault Tolerance][Phase2]" in mout and "[Fault Tolerance][Phase4]" in mout, mout
    assert "[Fault Tolerance][Phase3]" in relog and "[Fault Tolerance][Phase5]" in relog, relog
    assert "[Fault Tolerance][Phase5]" in out0, out0
    assert "generation 1 starts at iteration 40" in out0, out0     # killed at >= 50: last checkpoint is 40
    assert "rollbacks ordered for [2]" in mout, mout
    # result of the uninterrupted computation: w = 1e-3 * sum_it (it + 1) * (1 + 2)
    expect = 1e-3 * 3 * sum(range(1, 121))
    for i in (0, 2):
        w = torch.load(out / f"final_{i}.pt")
        assert torch.allclose(w, torch.full((8,), expect, dtype=torch.float64), rtol=0, atol=1e-9), (i, w)


def test_scale_out_new_node_joins_next_generation(tmp_path):
    """Elastic scale-out (reference lr_example.cpp:396-409, master_thread.cpp:40-43, mailbox.cpp:197-219):
    a job of ONE worker node is running; a new process announces itself (scale file + kScale), the master
    adds it, publishes the membership of the next generation and broadcasts kScaleRollback; both nodes
    meet there as ranks 0 and 1 of 2 and resume from the last checkpoint."""
    import re
    base = 22000 + os.getpid() % 2000
    hostfile = tmp_path / "nodes"
    hostfile.write_text(f"0:localhost:{base}\n1:localhost:{base + 1}\n")
    out = tmp_path / "out"
    out.mkdir()
    scale_file = str(tmp_path / "scale")
    args = {"config_file": str(hostfile), "heartbeat_interval": 1, "store_port": base + 7, "out_dir": str(out),
            "num_iters": 120, "checkpoint_interval": 20, "iter_sleep": 0.2, "scale_file": scale_file}
    worker = os.path.join(ROOT, "tests", "ft_worker.py")
    env = dict(os.environ, MINIPS_LOG="0")
    master = subprocess.Popen(
        [sys.executable, "-m", "minips_b200.apps.master", f"--config_file={hostfile}", "--heartbeat_interval=1",
         f"--store_port={base + 7}", f"--scale_file={scale_file}"], cwd=ROOT, env=env, stdout=subprocess.PIPE,
        stderr=subprocess.STDOUT, text=True)
    time.sleep(1.0)
    procs = {0: subprocess.Popen([sys.executable, worker, json.dumps(dict(args, my_id=0))], cwd=ROOT, env=env,
                                 stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)}
    try:
        assert _wait_for(out / "progress_0", lambda s: int(s) >= 25, 60), "node 0 never reached iteration 25"
        procs[2] = subprocess.Popen([sys.executable, worker, json.dumps(dict(args, my_id=2, scale_port=base + 2))],
                                    cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        outs = {}
        for i, p in procs.items():
            outs[i], _ = p.communicate(timeout=120)
            assert p.returncode == 0, outs[i]
        mout, _ = master.communicate(timeout=60)
        assert master.returncode == 0, mout
    finally:
        for p in list(procs.values()) + [master]:
            if p.poll() is None:
                p.kill()
    assert "scale-outs for [2]" in mout, mout
    assert open(scale_file).read().strip() == f"2:localhost:{base + 2}"
    m0 = re.search(r"generation 1 starts at iteration (\d+) as rank 0 of 2", outs[0])
    m2 = re.search(r"generation 1 starts at iteration (\d+) as rank 1 of 2", outs[2])
    assert m0 and m2 and m0.group(1) == m2.group(1), (outs[0], outs[2])
    c = int(m0.group(1))
    assert c >= 20 and c % 20 == 0
    # one node up to the checkpoint, two nodes after it: w = 1e-3 (sum_{it<c} (it+1) + 3 sum_{it>=c} (it+1))
    expect = 1e-3 * (sum(range(1, c + 1)) + 3 * sum(range(c + 1, 121)))
    for i in (0, 2):
        w = torch.load(out / f"final_{i}.pt")
        assert torch.allclose(w, torch.full((8,), expect, dtype=torch.float64), rtol=0, atol=1e-9), (i, w, expect)
