#!/usr/bin/env bash
# Round-2 single-GPU session: the whole GPU suite on the refactored kernels, bench lines, autotuner decisions, and one
# ncu --set full capture per default kernel (each only after the plain run of the same command exited 0).
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() { name=$1; shift; echo "== $name"; timeout 600 "$@" > "gpurun_out/$name.log" 2>&1; echo "exit $?"; }
run t_all python -m pytest tests -m gpu -x -q --timeout 300 -p no:cacheprovider; tail -4 gpurun_out/t_all.log | cut -c1-400
run smoke python -c "import __graft_entry__ as g; g.smoke()"; tail -2 gpurun_out/smoke.log | cut -c1-300
run bench_base python bench.py --steps 100 --warmup 20; tail -1 gpurun_out/bench_base.log | cut -c1-400
rm -f gpurun_out/autotune.txt
DCNN_AUTOTUNE=1 DCNN_AUTOTUNE_LOG=gpurun_out/autotune.txt run bench_autotune python bench.py --steps 100 --warmup 20
tail -1 gpurun_out/bench_autotune.log | cut -c1-300
run bench_ref python bench.py --impl reference --steps 20 --warmup 5; tail -1 gpurun_out/bench_ref.log | cut -c1-300
run kernels python scripts/prof_kernels.py; head -34 gpurun_out/kernels.log | cut -c1-170
if [ "${SKIP_NCU:-0}" != "1" ]; then
  timeout 120 python scripts/ncu_step.py > gpurun_out/ncu_plain.log 2>&1; rc=$?; echo "ncu target plain exit $rc"
  if [ $rc -eq 0 ]; then
    for k in conv_tc3_kernel conv_tc2_kernel wgrad_tc_kernel bn_bwd_apply_kernel bn_bwd_reduce_kernel bn_apply_kernel softmax_ce_kernel multi_tensor_update_kernel; do
      timeout 150 ncu --set full --clock-control none --import-source on --profile-from-start off -k "regex:$k" -c 1 -f \
        -o "gpurun_out/ncu_$k" python scripts/ncu_step.py > "gpurun_out/ncu_$k.log" 2>&1; echo "ncu $k exit $?"
    done
  fi
fi
ls -la gpurun_out | head -50
