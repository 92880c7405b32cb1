#!/usr/bin/env bash
# Third 8-GPU session of round 2: world-8 DP equivalence on the final kernels, the headline line (dp8 + pipeline child),
# 1F1B warm-up A/B, microbatch sweep, all-reduce roofline.
#   gpurun --gpus 8 --timeout 1200 -- 'bash scripts/gpu_r2_multi_c.sh'
N=8
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
line() { grep '^{' "$1" | tail -1 | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    p = d.get('pipeline') or {}
    print({k: d.get(k) for k in ('value', 'ms_per_step', 'n_gpus', 'pipeline_bubble_fraction', 'final_loss')}, d['config']['parallelism'], 'mb', d['config']['microbatches'], 'e2e', d['e2e']['value'], 'pipeline', {k: p.get(k) for k in ('parallelism', 'value', 'ms_per_step', 'microbatches', 'pipeline_bubble_fraction', 'unavailable')})
" 2>/dev/null || tail -3 "$1" | cut -c1-300; }
b() { name=$1; shift; echo "== $name"; timeout ${TMO:-240} env "$@" > "gpurun_out/$name.log" 2>&1; echo "exit $?"; line "gpurun_out/$name.log"; }
echo "== dp test world 8"; timeout 400 python -m pytest tests/test_gpu_dp.py -m gpu -q --timeout 280 -p no:cacheprovider -k "default" > gpurun_out/t_dp_8.log 2>&1; echo "exit $?"; tail -3 gpurun_out/t_dp_8.log | cut -c1-300
echo "== single GPU"; CUDA_VISIBLE_DEVICES=0 timeout 120 python bench.py --steps 100 --warmup 20 > gpurun_out/bench_single.log 2>&1; echo "exit $?"; line gpurun_out/bench_single.log
TMO=500 b bench_dp8 $TR bench.py --gpus 8 --steps 100 --warmup 20
echo "== allreduce roofline"; timeout 200 $TR benchmarks/allreduce_roofline.py > gpurun_out/ar_8.log 2>&1; echo "exit $?"; tail -1 gpurun_out/ar_8.log | cut -c1-700
b bench_pp4xdp2_mb16_f1 DCNN_PIPE_WARMUP_FACTOR=1 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp4xdp2 --microbatches 16
b bench_pp4xdp2_mb32 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp4xdp2 --microbatches 32
b bench_pp4xdp2_mb8 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp4xdp2 --microbatches 8
b bench_pp8_r34 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp8 --model resnet34_tiny_imagenet --microbatches 16
b bench_pp8_r34_mb32 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp8 --model resnet34_tiny_imagenet --microbatches 32
