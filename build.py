#!/usr/bin/env python
"""In-tree build of the native parts of minips_b200.

    python build.py            # incremental build of everything
    python build.py --force    # rebuild
    python build.py --sass     # also dump SASS listings into profiles/sass/

Targets
  minips_b200/lib/libminips_kernels.so   hand-written sm_100a CUDA kernels, C ABI
                                         (include/minips_kernels.h), loaded with ctypes
  minips_b200/lib/_minips_core*.so       C++17 host runtime + pybind11 bindings
  build/bin/*                            C++ executables (apps, host unit tests)

The .so files are git-ignored but travel to the GPU box with the gpurun snapshot.
nvcc cross-compiles for sm_100a on a GPU-less box.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import subprocess
import sys
import sysconfig
from pathlib import Path

ROOT = Path(__file__).resolve().parent
BUILD = ROOT / "build"
LIBDIR = ROOT / "minips_b200" / "lib"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17", "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v",
    f"-I{ROOT / 'include'}", f"-I{ROOT / 'csrc' / 'kernels'}",
]
CXX_FLAGS = ["-O2", "-g", "-std=c++17", "-fPIC", "-Wall", "-Wno-sign-compare", "-pthread",
             f"-I{ROOT / 'include'}", f"-I{ROOT / 'csrc'}"]


def _newer(src_deps, out: Path) -> bool:
    if not out.exists():
        return True
    t = out.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in src_deps)


def _run(cmd, log: Path | None = None):
    p = subprocess.run([str(c) for c in cmd], capture_output=True, text=True)
    if log is not None:
        log.parent.mkdir(parents=True, exist_ok=True)
        log.write_text(p.stdout + p.stderr)
    if p.returncode != 0:
        sys.stderr.write(" ".join(str(c) for c in cmd) + "\n" + p.stdout + p.stderr)
        raise SystemExit(f"build failed: {cmd[0]} ({p.returncode})")
    return p


def build_kernels(force=False, verbose=True) -> Path:
    src_dir = ROOT / "csrc" / "kernels"
    srcs = sorted(src_dir.glob("*.cu"))
    hdrs = list(src_dir.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    obj_dir = BUILD / "kernels"
    obj_dir.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    jobs = []
    for s in srcs:
        o = obj_dir / (s.stem + ".o")
        if force or _newer([s, *hdrs], o):
            jobs.append((s, o))
    def compile_one(job):
        s, o = job
        _run([NVCC, *NVCC_FLAGS, "-c", s, "-o", o], log=obj_dir / (s.stem + ".ptxas.log"))
        return s.name
    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for name in ex.map(compile_one, jobs):
            if verbose:
                print(f"[nvcc] {name}")
    out = LIBDIR / "libminips_kernels.so"
    objs = [obj_dir / (s.stem + ".o") for s in srcs]
    if force or jobs or not out.exists():
        _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out, *objs,
              "-lcudart", "-lcuda"])
        if verbose:
            print(f"[link] {out.relative_to(ROOT)}")
    return out


def core_sources():
    d = ROOT / "csrc" / "core"
    return sorted(d.glob("**/*.cpp")) if d.exists() else []


def build_core(force=False, verbose=True):
    """C++ host runtime: static objects -> python module + executables."""
    srcs = core_sources()
    if not srcs:
        return None
    import pybind11
    hdrs = list((ROOT / "csrc").glob("**/*.hpp"))
    obj_dir = BUILD / "core"
    obj_dir.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    jobs = []
    objs = []
    for s in srcs:
        o = obj_dir / (str(s.relative_to(ROOT / "csrc" / "core")).replace("/", "_")[:-4] + ".o")
        objs.append(o)
        if force or _newer([s, *hdrs], o):
            jobs.append((s, o, []))
    bind_src = ROOT / "csrc" / "bindings" / "core_bindings.cpp"
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    mod = LIBDIR / f"_minips_core{ext}"
    bind_obj = obj_dir / "core_bindings.o"
    py_inc = [f"-I{pybind11.get_include()}", f"-I{sysconfig.get_paths()['include']}"]
    if bind_src.exists() and (force or _newer([bind_src, *hdrs], bind_obj)):
        jobs.append((bind_src, bind_obj, py_inc))
    def compile_one(job):
        s, o, extra = job
        _run([CXX, *CXX_FLAGS, *extra, "-c", s, "-o", o])
        return s.name
    with cf.ThreadPoolExecutor(max_workers=8) as ex:
        for name in ex.map(compile_one, jobs):
            if verbose:
                print(f"[g++] {name}")
    if bind_src.exists() and (force or jobs or not mod.exists()):
        # --exclude-libs: a toolchain that links libstdc++ statically would otherwise export its copy, and
        # torch's (system) libstdc++ interposes half of it at run time — crashes in the first ostringstream
        _run([CXX, "-shared", "-o", mod, bind_obj, *objs, "-pthread", "-Wl,--exclude-libs,ALL"])
        if verbose:
            print(f"[link] {mod.relative_to(ROOT)}")
    # executables
    bin_dir = BUILD / "bin"
    bin_dir.mkdir(parents=True, exist_ok=True)
    exe_srcs = []
    for d in (ROOT / "csrc" / "apps", ROOT / "csrc" / "tests"):
        if d.exists():
            exe_srcs.extend(sorted(d.glob("*.cpp")))
    if exe_srcs:
        for s in exe_srcs:
            exe = bin_dir / s.stem
            if force or jobs or _newer([s, *hdrs], exe):
                _run([CXX, *CXX_FLAGS, s, *objs, "-o", exe, "-pthread"])
                if verbose:
                    print(f"[link] {exe.relative_to(ROOT)}")
    return mod


def dump_sass():
    out_dir = ROOT / "profiles" / "sass"
    out_dir.mkdir(parents=True, exist_ok=True)
    lib = LIBDIR / "libminips_kernels.so"
    p = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True)
    (BUILD / "libminips_kernels.sass").write_text(p.stdout)
    return BUILD / "libminips_kernels.sass"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--sass", action="store_true")
    ap.add_argument("--only", choices=["kernels", "core"], default=None)
    ap.add_argument("--host-only", action="store_true", help="C++ host runtime only (CI boxes without nvcc)")
    a = ap.parse_args()
    if a.host_only:
        a.only = "core"
    if a.only in (None, "kernels"):
        build_kernels(a.force)
    if a.only in (None, "core"):
        build_core(a.force)
    if a.sass:
        print(dump_sass())


if __name__ == "__main__":
    main()
