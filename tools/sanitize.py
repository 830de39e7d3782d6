#!/usr/bin/env python3
"""Race detection / sanitizer targets (the reference has none).

Host runtime (no GPU needed):
    python tools/sanitize.py host --tool thread      # TSan build of csrc/core + the C++ test suite
    python tools/sanitize.py host --tool address     # ASan + UBSan build, same suite

Device kernels (needs a B200; run through gpurun, ONE tool per call):
    python tools/sanitize.py device --tool memcheck  [--tests tests/test_gpu_push_pull.py ...]
    python tools/sanitize.py device --tool racecheck
    python tools/sanitize.py device --tool synccheck
    python tools/sanitize.py device --tool initcheck

The device target runs the selected GPU tests (default: the push/pull, sparse-utils and sorted-accumulate
tests, small and quick) in one pytest process under compute-sanitizer.  Cross-GPU flags are written with
st.relaxed.sys after a fence and read with ld.acquire/relaxed.sys (csrc/kernels/common.cuh), which
racecheck does not model: it checks shared-memory hazards inside the kernels.
"""
import argparse
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_DEVICE_TESTS = [
    "tests/test_gpu_push_pull.py",
    "tests/test_gpu_sparse_utils.py",
    "tests/test_gpu_kmeans.py::test_sorted_accumulate_matches_oracle",
]


def host(tool: str, filt: str) -> int:
    # the distribution compiler ships libtsan / libasan; a toolchain named by $CXX may not
    cxx = os.environ.get("MINIPS_SAN_CXX") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else
                                                os.environ.get("CXX", "g++"))
    san = {"thread": ["-fsanitize=thread"], "address": ["-fsanitize=address,undefined",
        "-fno-omit-frame-pointer"]}[tool]
    out_dir = os.path.join(ROOT, "build", f"san_{tool}")
    os.makedirs(out_dir, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "core", "**", "*.cpp"), recursive=True))
    srcs.append(os.path.join(ROOT, "csrc", "tests", "core_tests.cpp"))
    exe = os.path.join(out_dir, "core_tests")
    cmd = [cxx, "-std=c++17", "-O1", "-g", *san, f"-I{os.path.join(ROOT, 'csrc')}",
        f"-I{os.path.join(ROOT, 'csrc', 'core')}",
           f"-I{os.path.join(ROOT, 'include')}", *srcs, "-o", exe, "-pthread"]
    print("building", exe, flush=True)
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode:
        sys.stderr.write(p.stderr[-4000:])
        return p.returncode
    env = dict(os.environ)
    env.setdefault("TSAN_OPTIONS", "halt_on_error=0 second_deadlock_stack=1")
    env.setdefault("ASAN_OPTIONS", "detect_leaks=0")
    args = [exe] + ([filt] if filt else [])
    p = subprocess.run(args, env=env, capture_output=True, text=True, cwd=out_dir)
    log = os.path.join(out_dir, "report.txt")
    with open(log, "w") as f:
        f.write(p.stdout)
        f.write(p.stderr)
    warnings = p.stderr.count("WARNING: ThreadSanitizer") + p.stderr.count("ERROR: AddressSanitizer") + \
        p.stderr.count("runtime error:")
    print(p.stdout[-1500:])
    print(f"{tool} sanitizer: exit {p.returncode}, {warnings} report(s); full log {log}")
    return 1 if (p.returncode or warnings) else 0


def device(tool: str, tests) -> int:
    cmd = ["compute-sanitizer", "--tool", tool, "--error-exitcode", "86", "--target-processes", "application-only",
           sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", *tests]
    print(" ".join(cmd), flush=True)
    return subprocess.run(cmd, cwd=ROOT).returncode


def main() -> int:
    ap = argparse.ArgumentParser()
    sub = ap.add_subparsers(dest="what", required=True)
    h = sub.add_parser("host")
    h.add_argument("--tool", choices=["thread", "address"], default="thread")
    h.add_argument("--filter", default="", help="substring of the C++ test names to run")
    d = sub.add_parser("device")
    d.add_argument("--tool", choices=["memcheck", "racecheck", "synccheck", "initcheck"], default="memcheck")
    d.add_argument("--tests", nargs="*", default=DEFAULT_DEVICE_TESTS)
    a = ap.parse_args()
    if a.what == "host":
        return host(a.tool, a.filter)
    return device(a.tool, a.tests)


if __name__ == "__main__":
    sys.exit(main())
