mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 45 python scripts/ncu_conv.py > gpurun_out/conv_times.log 2>&1; rc=$?; echo "timing exit $rc"; cat gpurun_out/conv_times.log | cut -c1-200
if [ $rc -eq 0 ]; then
  NCU=1 timeout 60 ncu --set full --clock-control none --import-source on -k 'regex:conv_tc2|wgrad_tc' -c 3 -f -o gpurun_out/prof_conv_tc2 python scripts/ncu_conv.py > gpurun_out/ncu3.log 2>&1; echo "ncu exit $?"; tail -5 gpurun_out/ncu3.log | cut -c1-200
fi
