#!/usr/bin/env python3
"""BasicExample launcher (reference scripts/basic_example.py).  scripts/basic_example.py [local | gpu [N] | kill]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from launch_utils import ROOT, launch_util  # noqa: E402

hostfile = os.environ.get("MINIPS_HOSTFILE", os.path.join(ROOT, "config", "localnodes"))
if __name__ == "__main__":
    sys.exit(launch_util("basic_example", hostfile, {}, sys.argv, gpu_module="minips_b200.apps.basic_example") or 0)
