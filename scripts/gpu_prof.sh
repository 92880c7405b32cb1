mkdir -p gpurun_out
echo "== reference arm"; timeout 900 python bench.py --impl reference --steps 10 --warmup 5 > gpurun_out/bench_ref.log 2>&1; echo "exit $?"; tail -2 gpurun_out/bench_ref.log | cut -c1-700
echo "== plain"; timeout 300 python scripts/prof_step.py 256 3 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python scripts/prof_step.py 256 3 > gpurun_out/ncu.log 2>&1; echo "exit $?"
tail -3 gpurun_out/prof_plain.log; wc -l gpurun_out/launches.csv
