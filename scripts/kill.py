#!/usr/bin/env python3
"""killall of the app binaries on every host of a hostfile (reference scripts/kill.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from launch_utils import ROOT, kill_nodes  # noqa: E402

if __name__ == "__main__":
    hf = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "config", "localnodes")
    for prog in ("lr_example", "kmeans", "basic_example"):
        kill_nodes(prog, hf)
