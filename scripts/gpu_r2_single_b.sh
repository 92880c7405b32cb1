#!/usr/bin/env bash
# Round-2 single-GPU session B: full GPU suite, BASELINE.json single-GPU configs on both arms (>= 20 warm-up / 100 timed
# steps), per-op comparison with cuDNN, the product trainers, and one compute-sanitizer tool over the kernel tests.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() { name=$1; shift; echo "== $name"; timeout ${TMO:-300} "$@" > "gpurun_out/$name.log" 2>&1; echo "exit $?"; }
TMO=900 run t_all python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider; tail -8 gpurun_out/t_all.log | cut -c1-300
for m in resnet50_tiny_imagenet resnet9_cifar10 resnet18_tiny_imagenet resnet34_tiny_imagenet; do
  run cfg_${m}_ours python bench.py --model $m --batch 256 --steps 100 --warmup 20; tail -1 gpurun_out/cfg_${m}_ours.log | cut -c1-330
  run cfg_${m}_ref python bench.py --impl reference --model $m --batch 256 --steps 100 --warmup 20; tail -1 gpurun_out/cfg_${m}_ref.log | cut -c1-330
done
run conv_vs_cudnn python benchmarks/conv_vs_cudnn.py; tail -24 gpurun_out/conv_vs_cudnn.log | cut -c1-200
run kernels python scripts/prof_kernels.py; head -30 gpurun_out/kernels.log | cut -c1-170
if [ -n "${SANITIZER:-}" ]; then
  echo "== compute-sanitizer $SANITIZER"; bash scripts/compute_sanitizer.sh $SANITIZER "${SAN_EXPR:-conv or dense or batchnorm or pool or loss or optim}" > gpurun_out/sanitizer_${SANITIZER}_pytest.log 2>&1; echo "exit $?"
  tail -5 gpurun_out/sanitizer_${SANITIZER}_pytest.log | cut -c1-300; tail -6 gpurun_out/sanitizer_${SANITIZER}.log | cut -c1-300
fi
