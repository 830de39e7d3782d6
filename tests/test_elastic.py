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
    port = 21000 + os.getpid() % 2000
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
