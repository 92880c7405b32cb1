#!/bin/bash
# A/B of one environment switch on the flagship bench: gpu_ab_env.sh VAR v0 v1 ...   (results in gpurun_out/ab_VAR.log)
var=$1; shift
mkdir -p gpurun_out
: > gpurun_out/ab_$var.log
for v in "$@"; do
  echo "== $var=$v" >> gpurun_out/ab_$var.log
  env $var=$v BENCH_PIPELINE=0 python bench.py --steps 60 --warmup 15 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'])" >> gpurun_out/ab_$var.log
done
cat gpurun_out/ab_$var.log
