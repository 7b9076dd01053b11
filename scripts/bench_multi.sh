#!/bin/bash
# Every bench configuration (or those in $CONFIGS) on N GPUs of one box: bash scripts/bench_multi.sh N [tag]  -> gpurun_out/<tag>_bench_<config>_<N>gpu.json
N=${1:-2}
TAG=${2:-r02}
OUT=gpurun_out
mkdir -p $OUT
PORT=29540
for c in ${CONFIGS:-ws9_fused base_model post_chain train_dp}; do
  PORT=$((PORT+1))
  STEPS=10; [ "$c" = "train_dp" ] && STEPS=50
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N --config $c --steps $STEPS --warmup 5 > $OUT/${TAG}_bench_${c}_${N}gpu.json 2> $OUT/${TAG}_bench_${c}_${N}gpu.err
  echo "$c N=$N rc=$? $(tail -c 300 $OUT/${TAG}_bench_${c}_${N}gpu.json | head -c 0)"
  python - <<PY
import json
try:
    d=json.loads(open("$OUT/${TAG}_bench_${c}_${N}gpu.json").read().strip().splitlines()[-1])
    print("  ", d["n_gpus"], round(d["value"],1), d["unit"], "e2e", round(d["e2e"]["value"],1), d.get("extra",{}).get("allreduce_us"))
except Exception as e:
    print("  parse error", e)
PY
done
