#!/usr/bin/env bash
# First single-GPU call of the next hardware session: validates every opt-in feature written after the previous
# session's GPU minutes ran out, then measures what each is worth.  ~3-4 minutes of box time.
#   gpurun --timeout 420 -- 'bash scripts/gpu_next_session.sh'
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() { name=$1; shift; echo "== $name"; timeout 240 "$@" > "gpurun_out/$name.log" 2>&1; echo "exit $?"; }
DCNN_TEST_PENDING=1 run t_pending python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 240 -p no:cacheprovider \
    -k "bitmask or n256 or stem_packed or tc3 or batchnorm or narrow or autotuner"
tail -3 gpurun_out/t_pending.log | cut -c1-300
DCNN_CONV_2CTA=1 run check_2cta timeout 100 python scripts/check_conv_v3.py --pair; grep -v '^PASS' gpurun_out/check_2cta.log | cut -c1-200
echo "== t_layers"; DCNN_TEST_PENDING=1 timeout 900 python -m pytest tests/test_gpu_layers.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/t_layers.log 2>&1; echo "exit $?"
tail -3 gpurun_out/t_layers.log | cut -c1-300
run t_dp_first python -m pytest tests/test_gpu_dp.py -m gpu -q --timeout 300 -p no:cacheprovider   # needs >= 2 GPUs: skips on one
run t_all python -m pytest tests -m gpu -x -q --timeout 300 -p no:cacheprovider
tail -3 gpurun_out/t_all.log | cut -c1-300
for cfg in "base" "DCNN_BN_BITMASK=1" "DCNN_WGRAD_N256=1" "DCNN_STEM_TC=1" "DCNN_BN_FINALIZE_NARROW=1" "DCNN_CONV_2CTA=1" "DCNN_AUTOTUNE=1" "DCNN_BN_BITMASK=1 DCNN_WGRAD_N256=1 DCNN_STEM_TC=1 DCNN_BN_FINALIZE_NARROW=1"; do
  tag=$(echo "$cfg" | tr ' =' '__')
  if [ "$cfg" = "base" ]; then run "bench_$tag" python bench.py --steps 30 --warmup 5
  else run "bench_$tag" env $cfg python bench.py --steps 30 --warmup 5; fi
  tail -1 "gpurun_out/bench_$tag.log" | cut -c1-260
done
echo '== tma probe (native)'; [ -x scripts/native/tma_probe ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Icsrc scripts/native/tma_probe_main.cu csrc/core.cu -o scripts/native/tma_probe -lcuda
timeout 30 scripts/native/tma_probe gpurun_out/tma_probe2.txt; cat gpurun_out/tma_probe2.txt | cut -c1-200
for v in "--tf32 1 --channels-last 1 --graph 1" "--tf32 0 --channels-last 0 --graph 0"; do
  run "torch_baseline_$(echo $v | tr -d ' -')" python benchmarks/torch_baseline.py $v; tail -1 "gpurun_out/torch_baseline_$(echo $v | tr -d ' -').log" | cut -c1-400
done
run conv_vs_cudnn python benchmarks/conv_vs_cudnn.py; tail -25 gpurun_out/conv_vs_cudnn.log | cut -c1-200
DCNN_BN_BITMASK=1 DCNN_WGRAD_N256=1 DCNN_STEM_TC=1 DCNN_BN_FINALIZE_NARROW=1 run kernels_all python scripts/prof_kernels.py
head -30 gpurun_out/kernels_all.log | cut -c1-170
