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
cap lr_xchg_fp64 lr_train_xchg 1 python tools/profile_lr.py --ncu --dtype fp64
cap lr_xchg_fp32 lr_train_xchg 1 python tools/profile_lr.py --ncu --dtype fp32
cap kmeans_assign kmeans_assign_tc 1 python tools/profile_kmeans.py --ncu
ls -la $O
