#!/usr/bin/env bash
# The BASELINE.json configs other than the headline, through the same harness (bench.py --model ...), each under its own
# short timeout.  Usage on an N-GPU box:  bash scripts/gpu_baseline_configs.sh N      (N = 1, 2, 4 or 8)
N=${1:-1}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node"
run() { name=$1; shift; echo "== $name"; timeout 150 "$@" > "gpurun_out/cfg_$name.log" 2>&1; echo "exit $?"; tail -1 "gpurun_out/cfg_$name.log" | cut -c1-330; }
if [ "$N" -ge 1 ]; then
  run 2a_ours python bench.py --model resnet9_cifar10 --batch 256 --steps 30 --warmup 5
  run 2a_ref python bench.py --impl reference --model resnet9_cifar10 --batch 256 --steps 30 --warmup 5
  run rn18_ours python bench.py --model resnet18_tiny_imagenet --batch 256 --steps 30 --warmup 5
  run rn18_ref python bench.py --impl reference --model resnet18_tiny_imagenet --batch 256 --steps 30 --warmup 5
  run rn34_ours python bench.py --model resnet34_tiny_imagenet --batch 256 --steps 30 --warmup 5
  run rn34_ref python bench.py --impl reference --model resnet34_tiny_imagenet --batch 256 --steps 30 --warmup 5
fi
if [ "$N" -ge 2 ]; then   # config 3: ResNet-18, 2-stage sync pipeline with microbatching
  SCHEDULE=sync run 3_pp2_sync $TR 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --model resnet18_tiny_imagenet \
      --parallelism pp2 --microbatches 4 --steps 10 --warmup 3
fi
if [ "$N" -ge 8 ]; then
  run 2b_dp8 $TR 8 --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus 8 --model resnet9_cifar10 --batch 256 --steps 30 --warmup 5
  run 4_pp4xdp2 $TR 8 --master-addr 127.0.0.1 --master-port 29563 bench.py --gpus 8 --parallelism pp4xdp2 --microbatches 8 --steps 10 --warmup 3
  run 5_pp8 $TR 8 --master-addr 127.0.0.1 --master-port 29564 bench.py --gpus 8 --model resnet34_tiny_imagenet --parallelism pp8 \
      --microbatches 16 --steps 10 --warmup 3
fi
