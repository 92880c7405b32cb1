#!/usr/bin/env bash
# Second multi-GPU session of round 2: flag-in-data sync-BN, pipelined conv epilogue, reworked fused all-reduce.
#   gpurun --gpus 2 --timeout 900 -- 'bash scripts/gpu_r2_multi_b.sh 2'
N=${1:-2}
STEPS=${STEPS:-100}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
line() { grep '^{' "$1" | tail -1 | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    p = d.get('pipeline') or {}
    print({k: d.get(k) for k in ('value', 'ms_per_step', 'n_gpus', 'gpu_launches', 'pipeline_bubble_fraction', 'final_loss')}, d['config']['parallelism'], 'e2e', d['e2e']['value'], 'pipeline', {k: p.get(k) for k in ('parallelism', 'value', 'ms_per_step', 'microbatches', 'pipeline_bubble_fraction', 'unavailable')})
" 2>/dev/null || tail -3 "$1" | cut -c1-300; }
b() { name=$1; shift; echo "== $name"; timeout ${TMO:-200} env "$@" > "gpurun_out/$name.log" 2>&1; echo "exit $?"; line "gpurun_out/$name.log"; }
if [ "${SKIP_TESTS:-0}" != "1" ]; then
echo "== gpu tests"; timeout 900 python -m pytest ${TEST_SEL:-tests} -m gpu -q --timeout 280 -p no:cacheprovider ${TEST_K:+-k "$TEST_K"} > gpurun_out/t_all_$N.log 2>&1; echo "exit $?"; tail -12 gpurun_out/t_all_$N.log | cut -c1-400
fi
echo "== single GPU"; CUDA_VISIBLE_DEVICES=0 timeout 120 python bench.py --steps $STEPS --warmup 20 > gpurun_out/bench_single.log 2>&1; echo "exit $?"; line gpurun_out/bench_single.log
echo "== allreduce roofline"; timeout 200 $TR benchmarks/allreduce_roofline.py > gpurun_out/ar_$N.log 2>&1; echo "exit $?"; tail -1 gpurun_out/ar_$N.log | cut -c1-900
TMO=400 b bench_dp$N $TR bench.py --gpus $N --steps $STEPS --warmup 20
b bench_dp${N}_nosyncbn BENCH_SYNC_BN=0 BENCH_PIPELINE=0 $TR bench.py --gpus $N --steps $STEPS --warmup 20
b bench_dp${N}_buckets DCNN_DP_BUCKETS=4 BENCH_PIPELINE=0 $TR bench.py --gpus $N --steps $STEPS --warmup 20
if [ "$N" -eq 8 ]; then
  TMO=300 b bench_pp4xdp2_mb8 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp4xdp2 --microbatches 8
  b bench_pp8_r34 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp8 --model resnet34_tiny_imagenet --microbatches 16
  b bench_dp8_r9 BENCH_PIPELINE=0 $TR bench.py --gpus 8 --steps $STEPS --warmup 20 --model resnet9_cifar10
fi
