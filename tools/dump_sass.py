#!/usr/bin/env python3
"""Per-kernel SASS listings of libminips_kernels.so under profiles/sass/ (cuobjdump -sass, instruction
encodings stripped), plus the mnemonic histogram of every kernel (profiles/sass_mnemonics.txt).

    python tools/dump_sass.py            # the kernels of the named hot paths
    python tools/dump_sass.py --all      # every kernel (large)

What to look for (B200_PROFILING.md): UTCHMMA / UTCBAR = tcgen05.mma / commit, LDTM = tcgen05.ld,
UTMALDG = TMA loads, LDGMC / STG...MC... = multimem.ld_reduce / multimem.st, ...STRONG.SYS = peer
accesses, REDG = red.add at the owner."""
import argparse
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "minips_b200", "lib", "libminips_kernels.so")
HOT = [
    r"lr_train_xchg_kernel<8, double, false>", r"lr_train_xchg_kernel<8, float, false>",
    r"lr_train_xchg_kernel<8, double, true>", r"lr_train_dense_kernel<8, float>", r"lr_train_dense_kernel<8, double>",
    r"lr_step_sparse_kernel<8, float>", r"lr_eval_kernel<8, float>", r"kmeans_assign_tc_kernel<2>",
    r"kmeans_assign_tc_kernel<1>", r"kmeans_sorted_sum_kernel<128>", r"kmeans_delta_push_kernel",
    r"dense_lr_forward_kernel", r"dense_lr_grad_kernel", r"pull_dense_f32_kernel", r"push_dense_f32_kernel",
    r"pull_kernel<float>", r"push_kernel<float>", r"clock_kernel", r"csr_unpack_kernel", r"parse_lines_kernel",
    r"hash_add_kernel", r"hash_get_kernel", r"dense_lr_rows_kernel", r"dense_lr_stream_kernel",
    r"dense_lr_fused_kernel", r"lr_step_bucket_kernel<8, float>", r"lr_step_bucket_kernel<8, double>",
    r"fold_server_kernel",
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--all", action="store_true")
    a = ap.parse_args()
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    out_dir = os.path.join(ROOT, "profiles", "sass")
    os.makedirs(out_dir, exist_ok=True)
    for f in os.listdir(out_dir):
        if f.endswith(".sass"):
            os.unlink(os.path.join(out_dir, f))
    parts = re.split(r"\n\s*Function : ", txt)
    counts = {}
    written = []
    for part in parts[1:]:
        mangled, _, body = part.partition("\n")
        name = subprocess.run(["c++filt", mangled.strip()], capture_output=True, text=True).stdout.strip()
        short = re.sub(r"\(.*", "", name).replace("void ", "").replace("mps::", "")
        lines, hist = [], collections.Counter()
        for line in body.split("\n"):
            m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);\s*/\* 0x[0-9a-f]+ \*/", line)
            if m:
                ins = m.group(2).strip()
                lines.append(f"/*{m.group(1)}*/  {ins} ;")
                op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
                hist[op.split(".")[0]] += 1
        counts[short] = hist
        if a.all or any(h in name for h in HOT):
            fn = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_") + ".sass"
            with open(os.path.join(out_dir, fn), "w") as f:
                f.write(f"// {name}\n// {len(lines)} instructions; cuobjdump -sass of "
                        f"minips_b200/lib/libminips_kernels.so (sm_100a), encodings stripped\n")
                f.write("\n".join(lines) + "\n")
            written.append(fn)
    with open(os.path.join(ROOT, "profiles", "sass_mnemonics.txt"), "w") as f:
        f.write("# instruction mnemonics per kernel (tools/dump_sass.py); "
                "full listings of the hot kernels: profiles/sass/\n")
        keys = ("UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UBLKCP", "LDGMC", "STG", "REDG", "ATOMG", "LDG", "MEMBAR",
                "HMMA", "DFMA", "FFMA", "SHFL")
        for k in sorted(counts):
            h = counts[k]
            f.write(f"{k}: total={sum(h.values())} " + " ".join(f"{m}={h[m]}" for m in keys if h[m]) + "\n")
    print(f"wrote {len(written)} listings to profiles/sass/ and profiles/sass_mnemonics.txt")


if __name__ == "__main__":
    sys.exit(main())
