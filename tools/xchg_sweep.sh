#!/bin/bash
# Sweep of the LR exchange kernel's knobs on N GPUs: prints the in-kernel timeline and the per-step time.
N=${1:-2}
for b in ${BCASTS:-mc p2p}; do for d in ${DTYPES:-fp32 fp64}; do for r in ${REDUCES:-auto}; do
  MINIPS_XCHG_BCAST=$b MINIPS_XCHG_REDUCE=$r timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
    --master-addr 127.0.0.1 --master-port 29618 bench.py --gpus $N --steps 20 --warmup 5 --timeline --no-e2e \
    --no-secondary --dtype $d > gpurun_out/sw_${N}_${b}_${d}_${r}.json 2> gpurun_out/sw_${N}_${b}_${d}_${r}.err
  echo "bcast=$b reduce=$r $(grep timeline gpurun_out/sw_${N}_${b}_${d}_${r}.err)"
  grep -o '"ms_per_step": [0-9.e-]*' gpurun_out/sw_${N}_${b}_${d}_${r}.json | head -1
done; done; done
