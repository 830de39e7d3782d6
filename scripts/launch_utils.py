#!/usr/bin/env python3
"""Launch helpers (reference scripts/launch_utils.py:13-144, Python 2 + ssh): start one process
per hostfile line, kill them, relaunch one failed node from its checkpoint, add a scale-out node.

Two back ends:
  host  - the C++ actor runtime: `build/bin/<prog> --my_id=<id> --config_file=<hostfile> ...` per
          line; localhost lines are started as local subprocesses, other hosts through ssh.
  cuda  - the device parameter server: one `torchrun` with a rank per GPU of this box
          (`python -m minips_b200.apps.<module>`); the hostfile only fixes the node count.
"""
import os
import shlex
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LOCAL = ("localhost", "127.0.0.1")
ENV = {"MINIPS_VLOG": os.environ.get("MINIPS_VLOG", "1")}  # the reference sets GLOG_* here


def read_hostfile(path):
    nodes = []
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line and not line.startswith("#"):
                i, host, port = line.split(":")
                nodes.append((int(i), host, int(port)))
    return nodes


def params_to_args(params):
    return [f"--{k}={v}" for k, v in params.items()]


def _spawn(host, argv, env):
    if host in LOCAL:
        return subprocess.Popen(argv, env=dict(os.environ, **env))
    remote = " ".join(f"{k}={shlex.quote(v)}" for k, v in env.items()) + " " + " ".join(shlex.quote(a) for a in argv)
    return subprocess.Popen(["ssh", "-o", "StrictHostKeyChecking=no", host, remote])


def launch_nodes(prog, hostfile, params, env=None, wait=True):
    """Start <prog> for every node of the hostfile (reference launch_nodes :80-107)."""
    env = dict(ENV, **(env or {}))
    exe = prog if os.path.isabs(prog) else os.path.join(ROOT, "build", "bin", prog)
    procs = []
    for node_id, host, _ in read_hostfile(hostfile):
        argv = [exe, f"--my_id={node_id}", f"--config_file={os.path.abspath(hostfile)}"] + params_to_args(params)
        print("launch:", host, " ".join(argv))
        procs.append(_spawn(host, argv, env))
    if wait:
        return [p.wait() for p in procs]
    return procs


def launch_node(prog, hostfile, params, node_id, host, env=None):
    """Start ONE extra node (scale-out, reference launch_node :52-78)."""
    env = dict(ENV, **(env or {}))
    exe = prog if os.path.isabs(prog) else os.path.join(ROOT, "build", "bin", prog)
    argv = [exe, f"--my_id={node_id}", f"--config_file={os.path.abspath(hostfile)}"] + params_to_args(params)
    print("launch node:", host, " ".join(argv))
    return _spawn(host, argv, env)


def relaunch(prog, hostfile, params, failed_id, env=None):
    """Restart one failed node from its checkpoint (reference relaunch :13-50): same flags plus
    --use_weight_file=true.  This is what the master's `relaunch_cmd` points at."""
    params = dict(params, use_weight_file="true")
    for node_id, host, _ in read_hostfile(hostfile):
        if node_id == int(failed_id):
            return launch_node(prog, hostfile, params, node_id, host, env)
    raise SystemExit(f"node {failed_id} is not in {hostfile}")


def kill_nodes(prog, hostfile):
    """killall on every host of the hostfile (reference kill_nodes :110-123)."""
    name = os.path.basename(prog)
    for _, host, _ in read_hostfile(hostfile):
        cmd = ["killall", "-q", name] if host in LOCAL else ["ssh", host, f"killall -q {name}"]
        subprocess.call(cmd)


def launch_gpu(module, nproc, params, master_port=29533):
    """Device backend: one rank per GPU of this box."""
    argv = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
            "--master-addr", "127.0.0.1", "--master-port", str(master_port), "-m", module] + params_to_args(params)
    print("launch:", " ".join(argv))
    return subprocess.call(argv, cwd=ROOT, env=dict(os.environ, **ENV))


def launch_util(prog, hostfile, params, argv, gpu_module=None):
    """Mode dispatch shared by the per-app scripts (reference launch_util :136-144):
         <script> [local]            start every node of the hostfile
         <script> gpu [N]            device backend on N GPUs (default: all)
         <script> relaunch <id>      restart one node from its checkpoint
         <script> scale <id> <host> <port>
         <script> kill"""
    mode = argv[1] if len(argv) > 1 else "local"
    if mode == "kill":
        return kill_nodes(prog, hostfile)
    if mode in ("relaunch", "relocal"):
        return relaunch(prog, hostfile, params, argv[2]).wait()
    if mode == "scale":
        node_id, host, port = argv[2], argv[3], argv[4]
        p = dict(params, scale="true", scale_name=host, scale_port=port)
        return launch_node(prog, hostfile, p, node_id, host).wait()
    if mode == "gpu":
        import torch
        n = int(argv[2]) if len(argv) > 2 else max(1, torch.cuda.device_count())
        gp = {k: v for k, v in params.items() if k not in ("input", "assigner_master_port", "block_size")}
        return launch_gpu(gpu_module, n, gp)
    return launch_nodes(prog, hostfile, params)
