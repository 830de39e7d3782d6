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


def gpu_ft_commands(module, hostfile, params, store_port=29700, scale_file="/tmp/minips_scale"):
    """Command lines of a fault-tolerant device run: the master (node id 1) and one rank per other hostfile
    line, each bound to the GPU of its position (LOCAL_RANK).  Returns [(env, argv), ...], master first."""
    nodes = read_hostfile(hostfile)
    assert any(i == 1 for i, _, _ in nodes), "the fault-tolerant mode needs a node with id 1 (the master)"
    hb = params.get("heartbeat_interval", 1)
    hb = hb if int(hb) > 0 else 1
    hf = os.path.abspath(hostfile)
    relaunch = params.get("relaunch_cmd", "")
    master = [sys.executable, "-m", "minips_b200.apps.master", f"--config_file={hf}", f"--heartbeat_interval={hb}",
              f"--store_port={store_port}", f"--scale_file={scale_file}", f"--relaunch_cmd={relaunch}"]
    out = [({}, master)]
    rank_params = dict(params, heartbeat_interval=hb, store_port=store_port, scale_file=scale_file)
    rank_params.pop("relaunch_cmd", None)
    dev = 0
    for node_id, _, _ in nodes:
        if node_id == 1:
            continue
        out.append(({"LOCAL_RANK": str(dev)},
                    [sys.executable, "-m", module, f"--my_id={node_id}", f"--config_file={hf}"] +
                    params_to_args(rank_params)))
        dev += 1
    return out


def gpu_scale_command(module, hostfile, params, node_id, host, port, device, store_port=29700,
                      scale_file="/tmp/minips_scale"):
    """Command line of ONE new rank joining a running fault-tolerant device job (elastic scale-out)."""
    hb = params.get("heartbeat_interval", 1)
    hb = hb if int(hb) > 0 else 1
    p = dict(params, heartbeat_interval=hb, store_port=store_port, scale_file=scale_file, scale="true",
             scale_name=host, scale_port=port)
    p.pop("relaunch_cmd", None)
    return ({"LOCAL_RANK": str(device)},
            [sys.executable, "-m", module, f"--my_id={node_id}", f"--config_file={os.path.abspath(hostfile)}"] +
            params_to_args(p))


def gpu_relaunch_command(module, hostfile, params, failed_id, store_port=29700, scale_file="/tmp/minips_scale"):
    """Command line that restarts ONE rank of a gpuft job from the checkpoint (what the master's relaunch_cmd
    runs; reference relaunch :13-50 adds use_weight_file the same way)."""
    for env, argv in gpu_ft_commands(module, hostfile, params, store_port, scale_file)[1:]:
        if f"--my_id={int(failed_id)}" in argv:
            return env, argv + ["--use_weight_file=true", "--relaunched=true"]
    raise SystemExit(f"node {failed_id} is not a rank of {hostfile}")


def _run_all(cmds, delay_after_first=1.0):
    import time
    procs = []
    for k, (env, argv) in enumerate(cmds):
        print("launch:", " ".join(argv))
        procs.append(subprocess.Popen(argv, cwd=ROOT, env=dict(os.environ, **ENV, **env)))
        if k == 0:
            time.sleep(delay_after_first)     # the master hosts the rendezvous store
    return max(p.wait() for p in procs)


def launch_util(prog, hostfile, params, argv, gpu_module=None):
    """Mode dispatch shared by the per-app scripts (reference launch_util :136-144):
         <script> [local]            start every node of the hostfile
         <script> gpu [N]            device backend on N GPUs (default: all)
         <script> gpuft              device backend with master / heartbeats / rollback: master = node 1 of the
                                     hostfile, one rank per other line on GPU 0, 1, ...
         <script> gpuscale <id> <host> <port> <cuda index>   add a rank to a running gpuft job
         <script> gpurelaunch <id>   restart one rank of a gpuft job from its checkpoint (detached; this is
                                     the relaunch_cmd the master of a gpuft job runs)
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
    if mode in ("gpuft", "gpuscale", "gpurelaunch"):
        gp = {k: v for k, v in params.items() if k not in ("input", "assigner_master_port", "block_size")}
        gp["checkpoint_toggle"] = "true"      # nothing to roll back to otherwise
        gp["relaunch_cmd"] = f"{sys.executable} {os.path.abspath(argv[0])} gpurelaunch"
        if mode == "gpuft":
            return _run_all(gpu_ft_commands(gpu_module, hostfile, gp))
        if mode == "gpurelaunch":
            env, cmd = gpu_relaunch_command(gpu_module, hostfile, gp, argv[2])
            log = open(f"/tmp/minips_relaunch_{int(argv[2])}.log", "a")
            subprocess.Popen(cmd, cwd=ROOT, env=dict(os.environ, **ENV, **env), stdout=log, stderr=subprocess.STDOUT,
                             start_new_session=True)      # the master must not wait for it
            return 0
        return _run_all([gpu_scale_command(gpu_module, hostfile, gp, argv[2], argv[3], argv[4],
                                           argv[5] if len(argv) > 5 else 0)], 0.0)
    if mode == "gpu":
        import torch
        n = int(argv[2]) if len(argv) > 2 else max(1, torch.cuda.device_count())
        gp = {k: v for k, v in params.items() if k not in ("input", "assigner_master_port", "block_size")}
        return launch_gpu(gpu_module, n, gp)
    return launch_nodes(prog, hostfile, params)
