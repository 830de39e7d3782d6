#!/usr/bin/env python3
"""Style gate (reference scripts/lint.py + clang-format.py ran cpplint at 120 columns): here a
dependency-free check — line length <= 120 in C++/CUDA/Python sources, no tabs, no trailing
whitespace — plus `python -m compileall` of the package."""
import compileall
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXT = (".cpp", ".hpp", ".h", ".cu", ".cuh", ".py")
bad = 0
for top in ("csrc", "include", "minips_b200", "scripts", "tests", "baseline", "tools"):
    for d, _, files in os.walk(os.path.join(ROOT, top)):
        for f in files:
            if not f.endswith(EXT):
                continue
            p = os.path.join(d, f)
            for n, line in enumerate(open(p, errors="replace"), 1):
                s = line.rstrip("\n")
                if len(s) > 120 or "\t" in s or s != s.rstrip():
                    print(f"{os.path.relpath(p, ROOT)}:{n}: style ({len(s)} cols)")
                    bad += 1
ok = compileall.compile_dir(os.path.join(ROOT, "minips_b200"), quiet=1)
sys.exit(1 if bad or not ok else 0)
