#!/usr/bin/env python3
"""Build the UNMODIFIED reference (MiniPs) sources into baseline/_ref/bin/.

The reference is a CMake project whose dependencies are not in this image (libzmq headers, cppzmq,
glog, gflags, boost, libhdfs3, gtest) and cannot be fetched (no network).  Its sources are compiled
as they are, read-only from /root/reference, against the shim headers in baseline/ref_build/shims/
and linked with pyzmq's bundled libzmq.  Products: LRExample, KMeans, BasicExample.
"""
import concurrent.futures as cf
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("MINIPS_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "baseline", "_ref")
CXX = os.environ.get("CXX", "g++")


def find_libzmq():
    import zmq
    libs = os.path.join(os.path.dirname(os.path.dirname(zmq.__file__)), "pyzmq.libs")
    cands = sorted(glob.glob(os.path.join(libs, "libzmq-*.so*")))
    if not cands:
        raise SystemExit("pyzmq's bundled libzmq not found")
    return cands[0], libs


def main():
    os.makedirs(os.path.join(OUT, "obj"), exist_ok=True)
    os.makedirs(os.path.join(OUT, "bin"), exist_ok=True)
    libzmq, libdir = find_libzmq()
    flags = ["-std=c++11", os.environ.get("MINIPS_REF_OPT", "-O2"), "-w", "-pthread", "-DHAS_HADOOP", "-include",
        "cstdint", "-include", "cstring", "-include", "string",
             "-include", "functional", "-include", "memory", "-include", "algorithm", "-include", "fcntl.h",
             f"-I{os.path.join(HERE, 'shims')}", f"-I{REF}"]
    dirs = ["base", "comm", "driver", "server", "worker", "master", "io"]
    srcs, tests = [], []
    for d in dirs:
        for p in sorted(glob.glob(os.path.join(REF, d, "**", "*.cpp"), recursive=True)):
            if p.endswith("_test.cpp"):
                tests.append(p)
                continue
            srcs.append(p)
    apps = {"LRExample": os.path.join(REF, "apps/logistic_regression/lr_example.cpp"),
            "KMeans": os.path.join(REF, "apps/kmeans/kmeans.cpp"),
            "BasicExample": os.path.join(REF, "apps/basic/basic_example.cpp"),
            # our own microbenchmark program against the reference library (config 5 baseline)
            "RefPushPull": os.path.join(HERE, "ref_push_pull.cpp")}

    def obj_of(src):
        rel = os.path.relpath(src, REF) if src.startswith(REF) else "own_" + os.path.basename(src)
        return os.path.join(OUT, "obj", rel.replace("/", "_")[:-4] + ".o")

    shim_mtime = max(os.path.getmtime(p) for p in glob.glob(os.path.join(HERE, "shims", "**", "*.h*"), recursive=True))

    def compile_one(src):
        o = obj_of(src)
        if os.path.exists(o) and os.path.getmtime(o) > max(os.path.getmtime(src), shim_mtime):
            return src, 0, ""
        p = subprocess.run([CXX, *flags, "-c", src, "-o", o], capture_output=True, text=True)
        return src, p.returncode, p.stderr

    failed = False
    with cf.ThreadPoolExecutor(max_workers=8) as ex:
        tests.append(os.path.join(REF, "test", "test_main.cpp"))
        for src, rc, err in ex.map(compile_one, srcs + list(apps.values()) + tests):
            if rc:
                failed = True
                sys.stderr.write(f"--- {src}\n{err[:6000]}\n")
    if failed:
        raise SystemExit("reference build failed")
    objs = [obj_of(s) for s in srcs]
    for name, src in apps.items():
        exe = os.path.join(OUT, "bin", name)
        p = subprocess.run([CXX, "-pthread", obj_of(src), *objs, libzmq, f"-Wl,-rpath,{libdir}", "-o", exe],
                           capture_output=True, text=True)
        if p.returncode:
            sys.stderr.write(p.stderr[:6000])
            raise SystemExit(f"link of {name} failed")
        print("built", exe)
    # the reference's own unit tests (HuskyUnitTest in its CMake build) under the gtest shim
    exe = os.path.join(OUT, "bin", "RefUnitTest")
    p = subprocess.run([CXX, "-pthread", *[obj_of(t) for t in tests], *objs, libzmq, f"-Wl,-rpath,{libdir}", "-o", exe],
                       capture_output=True, text=True)
    if p.returncode:
        sys.stderr.write(p.stderr[:6000])
        raise SystemExit("link of RefUnitTest failed")
    print("built", exe)
    gen = os.path.join(OUT, "bin", "gen_libsvm")
    p = subprocess.run([CXX, "-std=c++17", "-O2", "-pthread", os.path.join(HERE, "gen_libsvm.cpp"), "-o", gen],
                       capture_output=True, text=True)
    if p.returncode:
        sys.stderr.write(p.stderr[:6000])
        raise SystemExit("build of gen_libsvm failed")
    print("built", gen)


if __name__ == "__main__":
    main()
