#!/usr/bin/env python3
"""KMeans launcher (reference scripts/kmeans.py:17-57).
   scripts/kmeans.py [local | gpu [N] | kill]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from launch_utils import ROOT, launch_util  # noqa: E402

hostfile = os.environ.get("MINIPS_HOSTFILE", os.path.join(ROOT, "config", "localnodes"))
data = os.environ.get("MINIPS_INPUT", "/tmp/minips_a9a.libsvm")
if not os.path.exists(data) and (len(sys.argv) < 2 or sys.argv[1] == "local"):
    sys.path.insert(0, ROOT)
    from minips_b200.ops.sparse import write_synthetic_libsvm
    write_synthetic_libsvm(data, 10000, 123, 14, seed=1)

params = {
    "input": data,
    "num_dims": 123,
    "K": 5,
    "batch_size": 6,
    "num_iters": 100,
    "alpha": 0.1,
    "kmeans_init_mode": "random",  # random | kmeans++ | kmeans_parallel
    "kModelType": "SSP",
    "kStaleness": 1,
    "report_interval": 10,
    "report_worker": 0,
    "num_workers_per_node": 1,
    "num_servers_per_node": 1,
    "num_local_load_thread": 2,
    "assigner_master_port": 19201,
    "block_size": 65536,
}
if len(sys.argv) > 1 and sys.argv[1] == "gpu":
    params.update({"num_dims": 128, "K": 1024, "batch_size": 1048576, "num_samples": 8000000, "kModelType": "BSP",
                   "alpha": 1.0})

if __name__ == "__main__":
    sys.exit(launch_util("kmeans", hostfile, params, sys.argv, gpu_module="minips_b200.apps.kmeans") or 0)
