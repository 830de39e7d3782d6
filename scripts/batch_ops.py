#!/usr/bin/env python3
"""Cluster helper (reference scripts/batch_ops.py: clone / pull / build the tree on a list of hosts).

    python scripts/batch_ops.py <hostfile> pull            # git pull in $MINIPS_HOME on every host
    python scripts/batch_ops.py <hostfile> build           # python build.py on every host
    python scripts/batch_ops.py <hostfile> run -- <cmd>    # any command, in $MINIPS_HOME

The hostfile is the usual `id:host:port` list (config/*); every distinct host is visited once over ssh
(`localhost` / 127.0.0.1 run directly).  MINIPS_HOME defaults to this checkout's path.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def hosts_of(path):
    seen = []
    for line in open(path):
        line = line.strip()
        if not line or line.startswith("#"):
            continue
        host = line.split(":")[1]
        if host not in seen:
            seen.append(host)
    return seen


def main(argv):
    if len(argv) < 3 or argv[2] not in ("pull", "build", "run"):
        print(__doc__)
        return 2
    home = os.environ.get("MINIPS_HOME", ROOT)
    if argv[2] == "pull":
        cmd = "git pull --ff-only"
    elif argv[2] == "build":
        cmd = f"{sys.executable} build.py"
    else:
        if "--" not in argv:
            print("run needs `-- <command>`")
            return 2
        cmd = " ".join(argv[argv.index("--") + 1:])
    rc = 0
    for host in hosts_of(argv[1]):
        full = f"cd {home} && {cmd}"
        local = host in ("localhost", "127.0.0.1")
        print(f"[{host}] {cmd}", flush=True)
        p = subprocess.run(["bash", "-lc", full] if local else
                           ["ssh", "-o", "StrictHostKeyChecking=no", host, full])
        rc = rc or p.returncode
    return rc


if __name__ == "__main__":
    sys.exit(main(sys.argv))
