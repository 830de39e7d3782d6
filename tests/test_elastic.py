"""Elastic membership of the device backend, host-side logic (CPU): ForceQuit shrink of the Engine
(reference comm/mailbox.cpp:159-171, driver/engine.cpp:33-36, lr_example.cpp:145-152) and the re-sharding
restore that lets a cluster that grew or shrank resume from a checkpoint (utils/checkpoint.py).  The
scale-out control flow is in tests/test_fault_tolerance.py, the device runs in tests/test_gpu_elastic.py."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_force_quit_shrinks_cluster():
    port = 31000 + os.getpid() % 1000
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3",
                        "--master-addr=127.0.0.1", f"--master-port={port}",
                        os.path.join(ROOT, "tests", "elastic_worker.py")],
                       cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "LEFT node=2" in p.stdout
    assert "STAYED node=0 rank=0 of 2" in p.stdout and "STAYED node=5 rank=1 of 2" in p.stdout


def test_restore_reshards_when_the_partition_changed(tmp_path):
    from minips_b200.utils import checkpoint as ck
    prefix = str(tmp_path / "ck_")
    rng = np.random.default_rng(0)
    full = rng.standard_normal(1000 * 2)             # 1000 keys, 2 values per key
    full[rng.random(full.size) < 0.3] = 0.0          # the dump format is sparse
    # dumped by 2 nodes, node 3 with two shards
    old = [(0, 0, 0, 400), (3, 0, 400, 700), (3, 1, 700, 1000)]
    for node, k, b, e in old:
        ck.write_sparse_line(ck._paths(prefix, node, k, "server_params"), full[2 * b:2 * e])
    layout = {"nodes": [0, 3], "vals_per_key": 2,
              "shards": [{"node": n, "k": k, "begin": b, "end": e} for n, k, b, e in old]}
    ck.write_layout(prefix, layout)
    assert ck.read_layout(prefix) == layout
    # restored by 3 nodes with one shard each, and by 1 node
    for parts in ([(0, 334), (334, 667), (667, 1000)], [(0, 1000)], [(399, 401), (699, 1000)]):
        cache = {}
        for b, e in parts:
            got = ck.read_resharded(prefix, layout, b, e, 2, np.float64, cache)
            np.testing.assert_array_equal(got, full[2 * b:2 * e])
    assert ck.read_layout(str(tmp_path / "nothing_")) is None


def test_launcher_builds_fault_tolerant_and_scale_out_commands(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import launch_utils as lu
    hostfile = tmp_path / "nodes"
    hostfile.write_text("0:localhost:9000\n1:localhost:9001\n2:localhost:9002\n")
    params = {"num_dims": 1000, "heartbeat_interval": -1, "relaunch_cmd": "x y"}
    cmds = lu.gpu_ft_commands("minips_b200.apps.lr_example", str(hostfile), params, store_port=29701,
                              scale_file="/tmp/s")
    assert len(cmds) == 3
    env0, master = cmds[0]
    assert "minips_b200.apps.master" in master and "--heartbeat_interval=1" in master and "--relaunch_cmd=x y" in master
    (e1, r1), (e2, r2) = cmds[1], cmds[2]
    assert e1 == {"LOCAL_RANK": "0"} and "--my_id=0" in r1 and "--store_port=29701" in r1
    assert e2 == {"LOCAL_RANK": "1"} and "--my_id=2" in r2 and not any(a.startswith("--relaunch_cmd") for a in r2)
    env, cmd = lu.gpu_scale_command("minips_b200.apps.lr_example", str(hostfile), params, 7, "hostb", 9100, 3)
    assert env == {"LOCAL_RANK": "3"}
    for a in ("--my_id=7", "--scale=true", "--scale_name=hostb", "--scale_port=9100", "--heartbeat_interval=1"):
        assert a in cmd, (a, cmd)
    env, cmd = lu.gpu_relaunch_command("minips_b200.apps.lr_example", str(hostfile), params, 2)
    assert env == {"LOCAL_RANK": "1"} and "--use_weight_file=true" in cmd and "--relaunched=true" in cmd
