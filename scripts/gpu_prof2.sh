mkdir -p gpurun_out
echo "== reference arm"; timeout 900 python bench.py --impl reference --steps 10 --warmup 5 > gpurun_out/bench_ref.log 2>&1; echo "exit $?"; tail -2 gpurun_out/bench_ref.log | cut -c1-900
echo "== ncu full conv_tc"; timeout 300 python scripts/prof_step.py 256 2 > gpurun_out/prof_plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 100 -c 4 -o gpurun_out/prof_conv_v1 python scripts/prof_step.py 256 2 > gpurun_out/ncu2.log 2>&1; echo "exit $?"; tail -3 gpurun_out/ncu2.log
ls -la gpurun_out/
