"""`bench.py --impl reference`: run the UNMODIFIED reference (MiniPs) from baseline/_ref.

The reference is a CMake/C++11 project (no setup.py / pyproject.toml), so
`pip install --target baseline/_ref /root/reference` cannot install it ("Directory is not
installable"; recorded in DESIGN.md).  baseline/ref_build/ builds its unmodified sources
against shim headers for the libraries this image lacks (libzmq headers, glog, gflags, boost,
libhdfs3, gtest); when that build exists this module drives its LRExample binary, otherwise it
reports the reference as unavailable.
"""
from __future__ import annotations

import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "baseline", "_ref", "bin", "LRExample")


def run(args):
    if getattr(args, "app", "lr") == "kmeans":
        if int(os.environ.get("RANK", "0")) != 0:
            return None
        return {"impl": "reference",
                "unavailable": "the reference KMeans app cannot run K=1024, d=128: it stores the per-centre counts in "
                               "params[K][i], a row of num_dims entries (kmeans_helper.hpp:41, needs K <= num_dims), "
                               "and "
                               "its random_init re-seeds rand() with time(NULL) per draw (one centre per second); see "
                               "BASELINE.md section 3"}
    if not os.path.exists(REF_BIN):
        if int(os.environ.get("RANK", "0")) != 0:
            return None
        return {"impl": "reference",
                "unavailable": "reference is a CMake C++ project: pip install reports 'not installable' and no "
                               "shim build of its sources exists in baseline/_ref (zmq/glog/gflags/boost/libhdfs3 "
                               "are not in this image)"}
    from ref_driver import run_lr
    return run_lr(args, REF_BIN)
