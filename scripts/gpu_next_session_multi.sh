#!/usr/bin/env bash
# First multi-GPU call of the next hardware session (every command under its own short timeout -- a hang costs N x):
#   gpurun --gpus 4 --timeout 300 -- 'bash scripts/gpu_next_session_multi.sh 4'
N=${1:-2}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
# default (validated) all-reduce first, then the opt-in loads-first gather (DCNN_AR_HOISTED=1) through the same tests
echo "== dp tests"; timeout 200 python -m pytest tests/test_gpu_dp.py -m gpu -x -q --timeout 180 -p no:cacheprovider > gpurun_out/t_dp_$N.log 2>&1; echo "exit $?"; tail -3 gpurun_out/t_dp_$N.log | cut -c1-300
echo "== dp tests, hoisted gather"; DCNN_AR_HOISTED=1 timeout 200 python -m pytest tests/test_gpu_dp.py -m gpu -x -q --timeout 180 -p no:cacheprovider > gpurun_out/t_dp_hoisted_$N.log 2>&1; echo "exit $?"; tail -3 gpurun_out/t_dp_hoisted_$N.log | cut -c1-300
echo "== allreduce roofline, hoisted gather"; DCNN_AR_HOISTED=1 timeout 200 $TR benchmarks/allreduce_roofline.py > gpurun_out/ar_hoisted_$N.log 2>&1; echo "exit $?"; tail -1 gpurun_out/ar_hoisted_$N.log | cut -c1-400
echo "== allreduce roofline"; timeout 200 $TR benchmarks/allreduce_roofline.py > gpurun_out/ar_$N.log 2>&1; echo "exit $?"; tail -1 gpurun_out/ar_$N.log | cut -c1-400
echo "== dp$N"; timeout 90 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_dp$N.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_dp$N.log | cut -c1-300
echo "== pp$N"; timeout 90 $TR bench.py --gpus $N --steps 10 --warmup 3 --parallelism pp$N --microbatches $((2*N)) > gpurun_out/bench_pp$N.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_pp$N.log | cut -c1-300
echo "== pp$N measured partition"; BENCH_PARTITION=measured timeout 120 $TR bench.py --gpus $N --steps 10 --warmup 3 --parallelism pp$N --microbatches $((2*N)) > gpurun_out/bench_pp${N}_measured.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_pp${N}_measured.log | cut -c1-300
if [ "$N" -ge 4 ]; then
  echo "== pp2xdp$((N/2))"; timeout 90 $TR bench.py --gpus $N --steps 10 --warmup 3 --parallelism pp2xdp$((N/2)) --microbatches 4 > gpurun_out/bench_pp2xdp.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_pp2xdp.log | cut -c1-300
fi
# message-driven pipeline (coordinator rank + N-1 stage ranks) on NCCL: the reference's deployment shape
echo "== message-driven semi-async pipeline, $N ranks"
SYNTHETIC=1 DEVICE_TYPE=GPU EPOCHS=1 BATCH_SIZE=128 NUM_MICROBATCHES=4 SYNTHETIC_TRAIN_SAMPLES=1024 SYNTHETIC_TEST_SAMPLES=256 \
  CUDA_MODULE_LOADING=EAGER STEP_TIMEOUT_S=60 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
  --master-addr 127.0.0.1 --master-port 29544 examples/semi_async_pipeline_coordinator.py > gpurun_out/msg_pipeline_$N.log 2>&1
echo "exit $?"; grep -E "Epoch|Validation|Stopped|images|Error|error" gpurun_out/msg_pipeline_$N.log | tail -8 | cut -c1-200
echo "== pp$N with per-op CUDA graphs"; DCNN_PIPE_GRAPH=1 timeout 120 $TR bench.py --gpus $N --steps 10 --warmup 4 --parallelism pp$N --microbatches $((2*N)) > gpurun_out/bench_pp${N}_graph.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_pp${N}_graph.log | cut -c1-300
# public training loops on GPUs (synthetic data): data-parallel trainer and SPMD pipeline trainer
COMMON="SYNTHETIC=1 DEVICE_TYPE=GPU EPOCHS=1 BATCH_SIZE=128 SYNTHETIC_TRAIN_SAMPLES=$((1024*N)) SYNTHETIC_TEST_SAMPLES=$((256*N)) PROGRESS_PRINT_INTERVAL=1000 STEP_TIMEOUT_S=60"
echo "== data_parallel_trainer x$N"; env $COMMON timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 examples/data_parallel_trainer.py --model resnet50_tiny_imagenet > gpurun_out/dp_trainer_$N.log 2>&1; echo "exit $?"; grep -E "Data parallel|Training   -|Validation -|Final|Error" gpurun_out/dp_trainer_$N.log | tail -5 | cut -c1-200
echo "== spmd_pipeline_trainer pp$N"; env $COMMON NUM_MICROBATCHES=4 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 examples/spmd_pipeline_trainer.py --model resnet50_tiny_imagenet --num-stages $N > gpurun_out/spmd_trainer_$N.log 2>&1; echo "exit $?"; grep -E "Pipeline:|Training   -|Validation -|Final|Error" gpurun_out/spmd_trainer_$N.log | tail -5 | cut -c1-200
echo "== dp$N with bucketed overlapped all-reduce"; DCNN_DP_BUCKETS=4 timeout 90 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_dp${N}_buckets.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_dp${N}_buckets.log | cut -c1-300
echo "== dp tests with buckets"; DCNN_DP_BUCKETS=3 timeout 200 python -m pytest tests/test_gpu_dp.py -m gpu -x -q --timeout 180 -p no:cacheprovider > gpurun_out/t_dp_buckets_$N.log 2>&1; echo "exit $?"; tail -3 gpurun_out/t_dp_buckets_$N.log | cut -c1-300
echo "== NVLS all-reduce (multimem)"; DCNN_DP_NVLS=1 timeout 90 $TR benchmarks/allreduce_roofline.py > gpurun_out/ar_nvls_$N.log 2>&1; echo "exit $?"; tail -2 gpurun_out/ar_nvls_$N.log | cut -c1-400
DCNN_DP_NVLS=1 timeout 200 python -m pytest tests/test_gpu_dp.py -m gpu -x -q --timeout 180 -p no:cacheprovider > gpurun_out/t_dp_nvls_$N.log 2>&1; echo "exit $?"; tail -3 gpurun_out/t_dp_nvls_$N.log | cut -c1-300
echo "== gpu pipeline tests"; DCNN_TEST_PENDING=1 timeout 400 python -m pytest tests/test_gpu_pipeline.py -m gpu -x -q --timeout 300 -p no:cacheprovider > gpurun_out/t_gpu_pipeline_$N.log 2>&1; echo "exit $?"; tail -3 gpurun_out/t_gpu_pipeline_$N.log | cut -c1-300
