#!/usr/bin/env python3
"""Format gate (reference scripts/clang-format.py -o check, run by its CI before the build).

With clang-format on PATH (or $CLANG_FORMAT): `clang-format --dry-run -Werror` over csrc/ and include/
against the repo's .clang-format; `-o replace` rewrites the files instead, like the reference's script.
Without it (this image has none) the structural part of the gate still runs on every source file:
LF line endings, a final newline, no tabs, no trailing blanks, at most 120 columns, header guards or
`#pragma once` in every header, and no `using namespace` at header scope.
"""
import argparse
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP_EXT = (".cpp", ".hpp", ".h", ".cu", ".cuh")
ALL_EXT = CPP_EXT + (".py",)
TOPS = ("csrc", "include", "minips_b200", "scripts", "tests", "baseline", "tools")
IGNORED = ("_ref", "__pycache__", "shims")


def sources(exts):
    for top in TOPS:
        for d, dirs, files in os.walk(os.path.join(ROOT, top)):
            dirs[:] = [x for x in dirs if x not in IGNORED]
            for f in sorted(files):
                if f.endswith(exts):
                    yield os.path.join(d, f)


def structural() -> int:
    bad = 0

    def report(p, n, what):
        nonlocal bad
        bad += 1
        print(f"{os.path.relpath(p, ROOT)}:{n}: {what}")

    for p in sources(ALL_EXT):
        raw = open(p, "rb").read()
        if b"\r\n" in raw:
            report(p, 1, "CRLF line endings")
        if raw and not raw.endswith(b"\n"):
            report(p, raw.count(b"\n") + 1, "no newline at end of file")
        text = raw.decode(errors="replace")
        for n, line in enumerate(text.split("\n"), 1):
            if "\t" in line:
                report(p, n, "tab")
            if line != line.rstrip():
                report(p, n, "trailing whitespace")
            if len(line) > 120:
                report(p, n, f"{len(line)} columns")
        if p.endswith((".hpp", ".h", ".cuh")):
            if "#pragma once" not in text and not re.search(r"#ifndef\s+\w+\s*\n\s*#define\s+\w+", text):
                report(p, 1, "header without include guard / #pragma once")
            depth = 0
            for n, line in enumerate(text.split("\n"), 1):
                if depth == 0 and re.match(r"\s*using\s+namespace\s+\w+", line):
                    report(p, n, "`using namespace` at header scope")
                depth += line.count("{") - line.count("}")
    return bad


def clang(mode: str) -> int:
    exe = os.environ.get("CLANG_FORMAT") or shutil.which("clang-format")
    if not exe:
        print("clang-format not found: structural checks only", file=sys.stderr)
        return 0
    files = [p for p in sources(CPP_EXT) if os.sep + "csrc" + os.sep in p or os.sep + "include" + os.sep in p]
    if mode == "replace":
        return subprocess.call([exe, "-i", "-style=file"] + files, cwd=ROOT)
    return subprocess.call([exe, "--dry-run", "-Werror", "-style=file"] + files, cwd=ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-o", dest="mode", default="check", choices=["check", "replace"])
    ap.add_argument("--strict-clang", action="store_true", help="fail when clang-format reports differences")
    a = ap.parse_args()
    bad = structural()
    rc = clang(a.mode)
    if rc and not a.strict_clang:
        print("clang-format reported differences (advisory without --strict-clang)", file=sys.stderr)
        rc = 0
    sys.exit(1 if bad or rc else 0)


if __name__ == "__main__":
    main()
