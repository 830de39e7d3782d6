#!/usr/bin/env python3
"""Push/pull microbenchmark of the REFERENCE build (baseline of BASELINE.json config 5).

Launches baseline/_ref/bin/RefPushPull (baseline/ref_build/ref_push_pull.cpp: our program against
the unmodified MiniPs library, ZMQ over localhost) as N node processes, takes the slowest node per
case, and prints keys/s and GB/s (4-byte key + 8-byte value per key on the wire, one direction).

    python tools/ref_push_pull.py --nodes 2 4 8 [--reps 7] [--out profiles/push_pull_reference.txt]

Compare with tools/bench_push_pull.py (our device path) at the same number of GPUs.
"""
import argparse
import os
import re
import socket
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "baseline", "_ref", "bin", "RefPushPull")


def free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def run(n, reps, sizes, timeout):
    work = tempfile.mkdtemp(prefix="ref_pp_")
    cfg = os.path.join(work, "nodes")
    with open(cfg, "w") as f:
        for i, p in enumerate(free_ports(n)):
            f.write(f"{i}:localhost:{p}\n")
    procs = [subprocess.Popen([EXE, f"--my_id={i}", f"--config_file={cfg}", f"--reps={reps}", f"--sizes={sizes}"],
                              stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, cwd=work)
             for i in range(n)]
    cases = {}
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise SystemExit(f"reference push/pull run with {n} nodes timed out")
        for m in re.finditer(r"RESULT node=(\d+) mode=(\w+) n=(\d+) push_ms=([0-9.]+) pull_ms=([0-9.]+)", out):
            key = (m.group(2), int(m.group(3)))
            push, pull = float(m.group(4)), float(m.group(5))
            old = cases.get(key, (0.0, 0.0, 0))
            cases[key] = (max(old[0], push), max(old[1], pull), old[2] + 1)
    return cases


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, nargs="+", default=[2, 4, 8])
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--sizes", default="1000,10000,100000,1000000,10000000")
    ap.add_argument("--timeout", type=float, default=600)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    if not os.path.exists(EXE):
        raise SystemExit("build first: python baseline/ref_build/build_ref.py")
    lines = [f"# reference (MiniPs, ZMQ over localhost, CPU) push/pull; host cores: {os.cpu_count()}",
             "# per node: keys/s = n / median time, slowest node; GB/s counts 12 bytes per key "
             "(uint32 key + double value)",
             f"{'nodes':>5} {'mode':>7} {'keys':>9} {'push_ms':>10} {'push_Mkeys/s':>13} {'push_GB/s':>10} "
             f"{'pull_ms':>10} {'pull_Mkeys/s':>13} {'pull_GB/s':>10} {'agg_pull_Mkeys/s':>17}"]
    for n in args.nodes:
        cases = run(n, args.reps, args.sizes, args.timeout)
        for (mode, k), (push, pull, cnt) in sorted(cases.items(), key=lambda kv: (kv[0][0] != "spread", kv[0][1])):
            if cnt != n:
                continue
            lines.append(f"{n:>5} {mode:>7} {k:>9} {push:>10.3f} {k / push / 1e3:>13.2f} {k * 12 / push / 1e6:>10.3f} "
                         f"{pull:>10.3f} {k / pull / 1e3:>13.2f} {k * 12 / pull / 1e6:>10.3f} "
                         f"{n * k / pull / 1e3:>17.2f}")
    text = "\n".join(lines)
    print(text)
    if args.out:
        with open(args.out, "w") as f:
            f.write(text + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
