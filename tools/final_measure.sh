#!/bin/bash
# Everything that is measured on N GPUs for the record (profiles/round2/): correctness worker, headline
# bench with the in-kernel timeline (fp64 headline, fp32 + k-means secondary, end-to-end), config 4
# (100M-key sparse regime) under SSP(2) and ASP.   usage: tools/final_measure.sh N
N=${1:-1}
O=gpurun_out/final_$N
mkdir -p $O
run() {  # run <port> <script> <args...>
  local port=$1; shift
  if [ "$N" = "1" ]; then python "$@"; else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $port "$@"; fi
}
if [ "$N" != "1" ]; then
  MINIPS_SPIN_TIMEOUT_S=10 run 29617 tests/multi_gpu_worker.py > $O/multi_gpu_worker.log 2>&1
  tail -n 1 $O/multi_gpu_worker.log
fi
run 29618 bench.py --gpus $N --steps 20 --warmup 5 --timeline > $O/bench.json 2> $O/bench.err
grep timeline $O/bench.err | tee $O/timeline.txt
for m in "SSP --staleness 2" "ASP"; do
  tag=$(echo $m | cut -d' ' -f1)
  MINIPS_BUCKET_TIMELINE=1 run 29619 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --no-secondary --dtype fp32 \
    --features 100000000 --model $m --timeline > $O/config4_$tag.json 2> $O/config4_$tag.err
  grep timeline $O/config4_$tag.err | tee -a $O/timeline.txt
done
run 29620 tools/h2d_ceiling.py 2>/dev/null | grep "h2d ceiling" | tee $O/h2d_ceiling.txt
python - <<PY
import json
for f in ("bench", "config4_SSP", "config4_ASP"):
    try:
        r = json.loads(open("$O/%s.json" % f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    line = "%s: %s %.3f G samples/s, %.2f us/step, %s" % (f, r["dtype"], r["value"] / 1e9, r["ms_per_step"] * 1e3,
                                                            r["config"].get("protocol"))
    if "e2e" in r:
        line += ", e2e %.3f G (%.1f us/step)" % (r["e2e"]["value"] / 1e9, r["e2e"]["ms_per_step"] * 1e3)
    for k, v in r.get("secondary", {}).items():
        if "value" in v:
            line += ", %s %.3f G" % (k, v["value"] / 1e9)
    print(line)
PY
