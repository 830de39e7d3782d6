#!/bin/bash
# ncu --set full captures of the top kernels (1 GPU): each only after the same command exited 0 without ncu.
# Reports land in gpurun_out/ncu/ (read here with `ncu -i ... --page raw --csv`, summaries in profiles/).
set -u
O=gpurun_out/ncu
mkdir -p $O
cap() {  # cap <name> <kernel regex> <count> <cmd...>
  local name=$1 rx=$2 cnt=$3; shift 3
  "$@" > $O/${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$rx -c $cnt -f -o $O/$name "$@" > $O/${name}_ncu.log 2>&1
  echo "$name: rc=$?"; tail -n 2 $O/${name}_ncu.log
}
# usage: tools/ncu_capture.sh [name ...]   (default: all)
want() { [ $# -eq 0 ] && return 0; local n=$1; shift; [ ${#SEL[@]} -eq 0 ] && return 0; for s in "${SEL[@]}"; do [ "$s" = "$n" ] && return 0; done; return 1; }
SEL=("$@")
want lr_xchg_fp64 && cap lr_xchg_fp64 lr_train_xchg 1 python tools/profile_lr.py --ncu --dtype fp64
want lr_xchg_fp32 && cap lr_xchg_fp32 lr_train_xchg 1 python tools/profile_lr.py --ncu --dtype fp32
want kmeans_assign && cap kmeans_assign kmeans_assign_tc 1 python tools/profile_kmeans.py --ncu
want dense_lr_rows && cap dense_lr_rows dense_lr_rows_kernel 1 python tools/profile_dense_lr.py --iters 2
want dense_lr_mma && MINIPS_DENSE_LR=mma cap dense_lr_mma dense_lr_fused_kernel 1 python tools/profile_dense_lr.py --iters 2
ls -la $O
