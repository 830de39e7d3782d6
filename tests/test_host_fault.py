"""Kill a node of the HOST runtime (the C++ lr_example, reference flow end to end): master (node 1) + nodes
0 and 2 train under SSP with checkpoints every 100 iterations; node 2 is SIGKILLed in mid-run; the master's
HeartBeatCheckThread declares it failed after 3 silent intervals and runs relaunch_cmd; the relaunched
process restores data / iteration / shard from the dumps (use_weight_file), its first heartbeat makes the
master broadcast kRollBack, node 0 rolls its servers and workers back, both meet on the fault barrier and
the run completes (reference master/heartbeat_check_thread.cpp:15-41, master_thread.cpp:19-43,
comm/mailbox.cpp:172-190, lr_example.cpp:265-288)."""
import json
import os
import signal
import subprocess
import sys
import time

from test_host_apps import ROOT, built, dataset  # noqa: F401 - fixtures


def _progress(path):
    try:
        return int(open(path).read().strip().split("\n")[-1].split("\t")[0])
    except (OSError, ValueError, IndexError):
        return 0


def test_host_kill_detect_relaunch_rollback(built, dataset, tmp_path):  # noqa: F811
    base = 29000 + os.getpid() % 480
    d = str(tmp_path)
    hostfile = tmp_path / "nodes"
    hostfile.write_text("".join(f"{i}:localhost:{base + i}\n" for i in range(3)))
    exe = os.path.join(built, "lr_example")
    common = [f"--config_file={hostfile}", f"--input={dataset}", "--num_dims=1000", "--kModelType=SSP",
              "--kStaleness=1", "--batch_size=50", "--num_iters=400", "--num_workers_per_node=2",
              "--heartbeat_interval=1", "--block_size=8192", f"--assigner_master_port={base + 5}",
              "--with_injected_straggler=1.0", "--checkpoint_toggle=true", "--checkpoint_interval=100",
              f"--checkpoint_file_prefix={d}/ck_", f"--checkpoint_raw_prefix={d}/ck_", "--init_dump=true",
              "--report_interval=10", f"--report_prefix={d}/rep"]
    launch = tmp_path / "launch.json"
    launch.write_text(json.dumps({"argv": [exe, "--my_id={id}"] + common, "log": f"{d}/relaunch_{{id}}.log",
                                  "append": ["--use_weight_file=true"]}))
    relaunch = f"--relaunch_cmd={sys.executable} {os.path.join(ROOT, 'scripts', 'relaunch_rank.py')} {launch}"
    logs = {i: open(f"{d}/out_{i}.log", "w") for i in range(3)}
    procs = {i: subprocess.Popen([exe, f"--my_id={i}"] + common + [relaunch], stdout=logs[i], stderr=subprocess.STDOUT)
             for i in range(3)}
    try:
        end = time.time() + 90
        while time.time() < end and _progress(f"{d}/rep_0") < 150:
            assert procs[2].poll() is None, open(f"{d}/out_2.log").read()[-3000:]
            time.sleep(0.2)
        assert _progress(f"{d}/rep_0") >= 150, "the run never reached iteration 150"
        procs[2].send_signal(signal.SIGKILL)
        procs[2].wait()
        for i in (0, 1):
            procs[i].wait(timeout=240)
    finally:
        for p in procs.values():
            if p.poll() is None:
                p.kill()
    out0, out1 = open(f"{d}/out_0.log").read(), open(f"{d}/out_1.log").read()
    relog = open(f"{d}/relaunch_2.log").read()
    assert procs[0].returncode == 0 and procs[1].returncode == 0, (out0[-3000:], out1[-3000:], relog[-3000:])
    assert "[Fault Tolerance][Phase2]" in out1 and "[Fault Tolerance][Phase4]" in out1, out1[-3000:]
    assert "[Fault Tolerance][Phase3]" in relog and "[Fault Tolerance][Phase5]" in relog, relog[-3000:]
    assert "Total training time" in out0 and "Total training time" in relog, (out0[-2000:], relog[-2000:])
