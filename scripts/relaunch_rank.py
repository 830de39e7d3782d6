#!/usr/bin/env python3
"""Relaunch hook of the master (reference scripts/launch_utils.py:13-50 `relaunch`: restart ONE failed
node with use_weight_file=True).  The master runs `relaunch_cmd + " " + <node id>`:

    python scripts/relaunch_rank.py <launch.json> <node id>

launch.json: {"argv": [... command line of a rank, with {id} where the node id goes ...],
              "env": {...extra environment...}, "log": "path with {id}", "device_of": {"<id>": <cuda index>}}.
The process is started detached (the master must not wait for it) with --use_weight_file=true --relaunched=true
appended."""
import json
import os
import subprocess
import sys


def main(argv):
    spec = json.load(open(argv[1]))
    node_id = argv[2]
    cmd = [a.replace("{id}", node_id) for a in spec["argv"]] + ["--use_weight_file=true", "--relaunched=true"]
    env = dict(os.environ)
    env.update({k: str(v).replace("{id}", node_id) for k, v in spec.get("env", {}).items()})
    dev = spec.get("device_of", {}).get(node_id)
    if dev is not None:
        env["LOCAL_RANK"] = str(dev)
    log = open(spec.get("log", "/tmp/minips_relaunch_{id}.log").replace("{id}", node_id), "a")
    subprocess.Popen(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, start_new_session=True)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
