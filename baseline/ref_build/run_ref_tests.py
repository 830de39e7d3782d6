#!/usr/bin/env python3
"""Run the reference's own unit tests (baseline/_ref/bin/RefUnitTest, built from its unmodified
*_test.cpp files under the gtest shim), one case per process.

One process per case because the reference's Mailbox never destroys its ZMQ context
(comm/mailbox.cpp:101-103), so a port released by zmq_close() in one case can still be bound when
the next case of the same process binds it again.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
EXE = os.path.join(ROOT, "baseline", "_ref", "bin", "RefUnitTest")


def main():
    if not os.path.exists(EXE):
        raise SystemExit("build first: python baseline/ref_build/build_ref.py")
    names = subprocess.run([EXE, "--gtest_list_tests"], capture_output=True, text=True, check=True).stdout.split()
    failed = []
    for name in names:
        try:
            p = subprocess.run([EXE, f"--gtest_filter={name}"], capture_output=True, text=True, timeout=120,
                               cwd="/tmp")
            ok = p.returncode == 0 and f"[       OK ] {name}" in p.stderr
        except subprocess.TimeoutExpired:
            ok = False
        print(("PASS " if ok else "FAIL ") + name, flush=True)
        if not ok:
            failed.append(name)
    print(f"{len(names) - len(failed)}/{len(names)} reference unit tests pass under the shims")
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
