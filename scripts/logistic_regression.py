#!/usr/bin/env python3
"""LRExample launcher (reference scripts/logistic_regression.py:17-93).
   scripts/logistic_regression.py [local | relaunch <id> | scale <id> <host> <port> | kill]      host runtime
   scripts/logistic_regression.py [gpu [N] | gpuft | gpuscale <id> <host> <port> <cuda> | gpurelaunch <id>]   device"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from launch_utils import ROOT, launch_util  # noqa: E402

hostfile = os.environ.get("MINIPS_HOSTFILE", os.path.join(ROOT, "config", "localnodes"))
data = os.environ.get("MINIPS_INPUT", "/tmp/minips_a9a.libsvm")
if not os.path.exists(data) and (len(sys.argv) < 2 or sys.argv[1] in ("local",)):
    sys.path.insert(0, ROOT)
    from minips_b200.ops.sparse import write_synthetic_libsvm
    write_synthetic_libsvm(data, 10000, 123, 14, seed=1)  # a9a-shaped stand-in: 123 features

params = {
    "input": data,
    "num_dims": 123,
    "kModelType": "SSP",      # ASP / SSP / BSP
    "kStorageType": "Vector",
    "kStaleness": 2,
    "batch_size": 100,
    "num_iters": 1000,
    "alpha": 0.1,
    "num_workers_per_node": 2,
    "num_servers_per_node": 1,
    "num_local_load_thread": 2,
    "with_injected_straggler": 0,
    "checkpoint_toggle": "true",
    "checkpoint_interval": 100,
    "checkpoint_file_prefix": "/tmp/minips_dump_",
    "checkpoint_raw_prefix": "/tmp/minips_dump_",
    "init_dump": "true",
    "heartbeat_interval": -1,
    "relaunch_cmd": f"{sys.executable} {os.path.abspath(__file__)} relaunch",
    "report_interval": -1,
    "assigner_master_port": 19201,
    "block_size": 65536,
}
if len(sys.argv) > 1 and sys.argv[1] in ("gpu", "gpuft", "gpuscale", "gpurelaunch"):
    params.update({"num_dims": 1000000, "batch_size": 65536, "num_samples": 2000000, "kModelType": "BSP",
                   "num_workers_per_node": 1, "checkpoint_toggle": "false"})

if __name__ == "__main__":
    sys.exit(launch_util("lr_example", hostfile, params, sys.argv, gpu_module="minips_b200.apps.lr_example") or 0)
