#!/usr/bin/env bash
# What the driver runs at round end, in one call: smoke(), the GPU test suite, both bench arms.
#   gpurun --timeout 900 -- 'bash scripts/gpu_round_end_check.sh'
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() { name=$1; shift; echo "== $name"; timeout 600 "$@" > "gpurun_out/$name.log" 2>&1; echo "exit $?"; }
run smoke python -c "import __graft_entry__ as g; g.smoke()"; tail -2 gpurun_out/smoke.log | cut -c1-300
run t_all python -m pytest tests -m gpu -x -q --timeout 300 -p no:cacheprovider; tail -4 gpurun_out/t_all.log | cut -c1-300
run bench_base python bench.py; tail -1 gpurun_out/bench_base.log | cut -c1-1200
run bench_ref python bench.py --impl reference; tail -1 gpurun_out/bench_ref.log | cut -c1-600
