#!/usr/bin/env bash
# Round-2 multi-GPU session (every command under its own short timeout -- a hang costs N x):
#   gpurun --gpus 2 --timeout 900 -- 'bash scripts/gpu_r2_multi.sh 2'
N=${1:-2}
STEPS=${STEPS:-50}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
line() { grep '^{' "$1" | tail -1 | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print({k: d.get(k) for k in ('value', 'ms_per_step', 'n_gpus', 'gpu_launches', 'pipeline_bubble_fraction', 'final_loss')}, d['config']['parallelism'], 'e2e', d['e2e']['value'])
" 2>/dev/null || tail -3 "$1" | cut -c1-300; }
b() { name=$1; shift; echo "== $name"; timeout 150 env "$@" > "gpurun_out/$name.log" 2>&1; echo "exit $?"; line "gpurun_out/$name.log"; }
if [ "${SKIP_TESTS:-0}" != "1" ]; then
echo "== all gpu tests"; timeout 900 python -m pytest tests -m gpu -q --timeout 280 -p no:cacheprovider > gpurun_out/t_all_$N.log 2>&1; echo "exit $?"; tail -15 gpurun_out/t_all_$N.log | cut -c1-400
fi
echo "== allreduce roofline"; timeout 200 $TR benchmarks/allreduce_roofline.py > gpurun_out/ar_$N.log 2>&1; echo "exit $?"; tail -1 gpurun_out/ar_$N.log | cut -c1-600
echo "== allreduce roofline NVLS"; DCNN_DP_NVLS=1 timeout 200 $TR benchmarks/allreduce_roofline.py > gpurun_out/ar_nvls_$N.log 2>&1; echo "exit $?"; tail -1 gpurun_out/ar_nvls_$N.log | cut -c1-600
b bench_dp$N $TR bench.py --gpus $N --steps $STEPS --warmup 10
b bench_dp${N}_nosyncbn BENCH_SYNC_BN=0 $TR bench.py --gpus $N --steps $STEPS --warmup 10
b bench_dp${N}_nccl BENCH_NCCL_BASELINE=1 $TR bench.py --gpus $N --steps $STEPS --warmup 10
b bench_dp${N}_buckets DCNN_DP_BUCKETS=4 $TR bench.py --gpus $N --steps $STEPS --warmup 10
b bench_dp${N}_nvls DCNN_DP_NVLS=1 $TR bench.py --gpus $N --steps $STEPS --warmup 10
if [ "$N" -eq 2 ]; then
  b bench_pp2_r18_sync SCHEDULE=sync $TR bench.py --gpus 2 --steps 20 --warmup 5 --parallelism pp2 --model resnet18_tiny_imagenet --microbatches 4
  b bench_pp2_r18_sync_graph SCHEDULE=sync DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 2 --steps 20 --warmup 5 --parallelism pp2 --model resnet18_tiny_imagenet --microbatches 4
  b bench_pp2_r50_graph DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 2 --steps 20 --warmup 5 --parallelism pp2 --microbatches 8
fi
if [ "$N" -ge 4 ]; then
  b bench_pp${N}_graph DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus $N --steps 20 --warmup 5 --parallelism pp$N --microbatches $((4*N))
fi
if [ "$N" -eq 8 ]; then
  for mb in 8 16; do
    b bench_pp4xdp2_mb$mb DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp4xdp2 --microbatches $mb
  done
  b bench_pp8_r34 DCNN_PIPE_GRAPH=1 BENCH_PARTITION=measured $TR bench.py --gpus 8 --steps 20 --warmup 5 --parallelism pp8 --model resnet34_tiny_imagenet --microbatches 16
fi
if [ "${SKIP_EXAMPLES:-0}" != "1" ]; then
COMMON="SYNTHETIC=1 DEVICE_TYPE=GPU EPOCHS=1 BATCH_SIZE=128 SYNTHETIC_TRAIN_SAMPLES=$((1024*N)) SYNTHETIC_TEST_SAMPLES=$((256*N)) PROGRESS_PRINT_INTERVAL=1000 STEP_TIMEOUT_S=60 CUDA_MODULE_LOADING=EAGER NUM_MICROBATCHES=4"
for ex in semi_async_pipeline_coordinator sync_pipeline_coordinator coordinator_tiny_imagenet; do
  echo "== $ex x$N"; env $COMMON timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 examples/$ex.py > gpurun_out/ex_${ex}_$N.log 2>&1
  echo "exit $?"; grep -E "Epoch|Validation|Stopped|images|Error|error|Traceback" gpurun_out/ex_${ex}_$N.log | tail -6 | cut -c1-200
done
echo "== data_parallel_trainer x$N"; env $COMMON timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 examples/data_parallel_trainer.py --model resnet50_tiny_imagenet > gpurun_out/ex_dp_trainer_$N.log 2>&1; echo "exit $?"; grep -E "Data parallel|Training   -|Validation -|Final|Error|Traceback" gpurun_out/ex_dp_trainer_$N.log | tail -5 | cut -c1-200
echo "== spmd_pipeline_trainer pp$N"; env $COMMON timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 examples/spmd_pipeline_trainer.py --model resnet50_tiny_imagenet --num-stages $N > gpurun_out/ex_spmd_trainer_$N.log 2>&1; echo "exit $?"; grep -E "Pipeline:|Training   -|Validation -|Final|Error|Traceback" gpurun_out/ex_spmd_trainer_$N.log | tail -5 | cut -c1-200
fi
