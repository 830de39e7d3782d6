"""Drive the UNMODIFIED reference LRExample (baseline/_ref/bin, built by baseline/ref_build) for
`bench.py --impl reference`.

Mapping of the bench contract onto the reference (a CPU parameter server):

* The reference does not use GPUs, so for every N it runs its 8-server-shard deployment on this
  box: 8 node processes (ids 0,2..8; id 1 is the reference's master/heartbeat process, see
  base/node_utils.cpp:17-31 — the LR app's no-master mode reads an uninitialised Node, so the
  master is always started), ZMQ over localhost.  The LR app starts exactly one server thread per
  node (lr_example.cpp:136 calls StartEverything() with its default), hence 8 nodes for 8 shards.
* BSP, Vector storage; global batch per step = N * --batch rows (the same as our arm at N GPUs),
  split evenly over 8 * workers_per_node worker threads so the CPU build can use the host's cores.
* data: the same synthetic shape (rows x features, nnz/row) written as a libsvm text file and
  read through the reference's own HDFS loader (hdfs shim on the local file system).
* time: the reference's own timer ("Total training time", lr_example.cpp:238-376) which brackets
  exactly num_iters iterations of Get -> gradient -> Add -> Clock per worker; max over workers.
  Loading, parsing and batch-key preparation are outside that timer, connections are set up
  before it.  The unmodified app has no separate warm-up phase, so W extra iterations are run
  inside the timer and the per-step time is total / (W + K).
"""
from __future__ import annotations

import os
import re
import subprocess
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NNZ_PER_ROW = 40
N_SHARDS = 8


def _free_ports(n):
    import socket
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def _dataset(gen, rows, features, nnz):
    d = os.environ.get("MINIPS_REF_DATA_DIR", os.path.join(tempfile.gettempdir(), "minips_ref_data"))
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, f"lr_{rows}x{features}x{nnz}.libsvm")
    if not os.path.exists(path):
        tmp = path + f".{os.getpid()}.tmp"
        subprocess.run([gen, tmp, str(rows), str(features), str(nnz), "1"], check=True)
        os.replace(tmp, path)
    return path


def _wait_for(paths, needle, procs, timeout):
    t0 = time.time()
    while time.time() - t0 < timeout:
        done = 0
        for p in paths:
            try:
                with open(p, "rb") as f:
                    if needle in f.read():
                        done += 1
            except OSError:
                pass
        if done == len(paths):
            return True
        if any(pr.poll() not in (None, 0) for pr in procs):
            return False
        time.sleep(0.25)
    return False


def run_lr(args, ref_bin):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None                       # rank 0 launches every reference process of the job
    n = max(1, int(os.environ.get("WORLD_SIZE", args.gpus)))
    n_req = n
    warm = max(3, args.warmup)
    # the reference's gradient loop scans the batch's whole key list once per row (lr_example.cpp:299-311),
    # so a step costs seconds at these batch sizes: bound the number of timed iterations per run
    cap = int(os.environ.get("MINIPS_REF_MAX_STEPS", 0)) or {1: 90, 2: 60}.get(n_req, 30)
    # and never reach iteration 100: worker 0 then calls table->CheckPoint() unconditionally
    # (lr_example.cpp:320-331), and with checkpoint_toggle=false the servers never answer it
    warm = min(warm, 10)
    steps = min(args.steps, cap, 100 - warm)
    rows = int(os.environ.get("MINIPS_REF_SAMPLES", args.samples))
    features = args.features
    gen = os.path.join(os.path.dirname(ref_bin), "gen_libsvm")
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 8)
    n_gpus = n
    n = int(os.environ.get("MINIPS_REF_NODES", N_SHARDS))        # reference node processes
    servers = 1
    # that scan makes the work per step fall with the number of workers the batch is split over (rows per
    # worker x keys per worker), so many worker threads per node is the reference's best setting even when
    # they outnumber the cores
    wpn = int(os.environ.get("MINIPS_REF_WORKERS_PER_NODE", 0)) or min(8 * n_gpus, 32)
    global_batch = n_gpus * args.batch
    while global_batch % (n * wpn):
        wpn -= 1
    batch = global_batch // (n * wpn)
    load_threads = max(1, min(4, cores // n))
    limit = float(os.environ.get("MINIPS_REF_TIMEOUT", 900))
    # memory: every reference worker thread deep-copies its node's samples (BatchDataSampler's constructor,
    # lib/batch_data_sampler.cpp:28-32: `data_ = data`), ~800 B per 40-nnz row, so the job needs about
    # rows x 800 B x (workers_per_node + 1).  Keep that under 45 % of the free host memory by loading fewer
    # rows of the same distribution (batches are random windows, so the step cost does not depend on it).
    rows_requested = rows
    try:
        with open("/proc/meminfo") as f:
            avail = next(int(l.split()[1]) * 1024 for l in f if l.startswith("MemAvailable"))
    except (OSError, StopIteration):
        avail = 64 << 30
    rows_fit = int(0.45 * avail / (800 * (wpn + 1)))
    if "MINIPS_REF_SAMPLES" not in os.environ and rows > rows_fit:
        rows = max(200_000, rows_fit // 100_000 * 100_000)

    t_gen = time.time()
    data = _dataset(gen, rows, features, NNZ_PER_ROW)
    t_gen = time.time() - t_gen

    work = tempfile.mkdtemp(prefix="minips_ref_run_")
    ids = [0] + list(range(2, n + 1))                 # worker/server nodes; id 1 is the master
    ports = _free_ports(n + 2)
    cfg = os.path.join(work, "nodes")
    with open(cfg, "w") as f:
        for i, node in enumerate([0, 1] + ids[1:]):
            f.write(f"{node}:localhost:{ports[i]}\n")
    hb = 20          # a node counts as failed after 3 intervals: leave room for 250+ busy threads per box
    common = [
        f"--config_file={cfg}", "--hdfs_namenode=localhost", "--hdfs_namenode_port=9000",
        f"--input=hdfs://{data}", f"--kModelType={args.model}", "--kStorageType=Vector",
        f"--num_dims={features}", f"--batch_size={batch}", f"--num_iters={steps + warm}",
        f"--kStaleness={args.staleness}", "--kSpeculation=0", f"--num_workers_per_node={wpn}",
        f"--num_servers_per_node={servers}", f"--num_local_load_thread={load_threads}",
        "--with_injected_straggler=0", f"--alpha={args.alpha}", f"--heartbeat_interval={hb}",
        "--scale=false", "--init_dump=false", "--checkpoint_toggle=false", "--report_interval=-1",
        f"--assigner_master_port={ports[n + 1]}", "--relaunch_cmd=true ",
    ]
    # the loader hands out "HDFS blocks" first come first served and a node that gets none leaves the
    # job (ForceQuit): make the blocks small enough that every loader thread sees many of them
    size = os.path.getsize(data)
    block = min(64 << 20, max(128 << 10, size // (n * load_threads * 64)))
    # equal-sized blocks: a tail block shorter than one line yields an empty record, on which the
    # reference's loader thread stops (lib/abstract_data_loader.hpp:41) — with no data if it was its first
    nblocks = -(-size // block)
    block = -(-size // nblocks)
    env = dict(os.environ, GLOG_minloglevel="0", MINIPS_REF_HDFS_BLOCK=str(block))
    procs, errs = [], []

    def launch(node):
        err = os.path.join(work, f"node{node}.err")
        p = subprocess.Popen([ref_bin, f"--my_id={node}", *common], env=env, cwd=work,
                             stdout=subprocess.DEVNULL, stderr=open(err, "wb"))
        procs.append(p)
        return err

    t_load = time.time()
    for node in ids:
        errs.append(launch(node))
    # the master counts a node as failed after 3 heartbeat intervals without a beat, and beats only
    # start after loading: start the master once every node has its data
    deadline = time.time() + limit
    ok = _wait_for(errs, b"Finished loading data", procs, timeout=limit)
    t_load = time.time() - t_load
    if ok:
        launch(1)
        while True:
            rcs = [p.poll() for p in procs[:n]]
            if all(rc is not None for rc in rcs):
                break
            if any(rc not in (None, 0) for rc in rcs) or time.time() > deadline:
                ok = False                   # a node died or the run is stuck: stop the others
                break
            time.sleep(0.2)
        try:
            procs[n].wait(timeout=10)        # the master leaves when every node has said goodbye
        except subprocess.TimeoutExpired:
            pass
    times, acc = [], []
    for e in errs:
        with open(e, "r", errors="replace") as f:
            txt = f.read()
        times += [float(m) for m in re.findall(r"Total training time: (\d+) ms", txt)]
        acc += [float(m) for m in re.findall(r"The accuracy is ([0-9.]+)", txt)]
    for p in procs:
        if p.poll() is None:
            p.kill()
            p.wait()
    rcs = [p.returncode for p in procs[:n]]
    if not ok or len(times) != n * wpn or any(rc != 0 for rc in rcs):
        tail = ""
        try:
            with open(errs[0], "r", errors="replace") as f:
                tail = f.read()[-300:].replace("\n", " | ")
        except OSError:
            pass
        return {"impl": "reference",
                "unavailable": (f"reference run failed (exit codes {rcs}, {len(times)}/{n * wpn} worker timings): "
                                f"{tail}")}
    total_ms = max(times)
    ms_per_step = total_ms / (steps + warm)
    value = global_batch / (ms_per_step / 1e3)
    return {
        "metric": "sparse_lr_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": n_gpus,
        "steps": steps, "steps_requested": args.steps, "warmup": warm, "ms_per_step": ms_per_step,
        "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "fp64 parameters (reference KVClientTable<double>), fp32 features",
        "data": "synthetic", "impl": "reference",
        "config": {
            "model": "sparse_logistic_regression", "samples": rows, "samples_requested": rows_requested,
            "features": features, "nnz_per_row": NNZ_PER_ROW,
            "server_shards": servers * n, "consistency": args.model, "staleness": args.staleness,
            "global_batch": global_batch, "seq_len": None,
            "parallelism": f"{n} reference node processes on this box x ({wpn} worker + {servers} server thread(s)) "
                           "+ master, ZMQ mailbox over localhost, CPU only (the reference has no GPU path)",
            "reference_nodes": n,
            "workers_per_node": wpn, "batch_per_worker": batch, "host_cores": cores,
            "timed_region": "the reference's own 'Total training time' (steady_clock around all W+K iterations of "
                            "Get/gradient/Add/Clock), max over workers; W warm-up iterations cannot be excluded in "
                            "the unmodified app, so ms_per_step = total / (W + K)",
            "build": "unmodified /root/reference sources, g++ -O2, shim headers for zmq.hpp/glog/gflags/boost/"
                     "libhdfs3 (baseline/ref_build), pyzmq's bundled libzmq",
            "accuracy": acc[:1], "datagen_s": round(t_gen, 1), "load_s": round(t_load, 1),
        },
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "CPU implementation: no device copies; same number as value"},
        "gpu_launches": 0,
    }
