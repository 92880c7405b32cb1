mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
echo "== bn tests"; timeout 60 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 60 -k "batchnorm" -p no:cacheprovider > gpurun_out/t_bn3.log 2>&1; echo "exit $?"; tail -4 gpurun_out/t_bn3.log | cut -c1-300
echo "== bench default"; timeout 60 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_final.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_final.log | cut -c1-330
echo "== v3 check"; DCNN_CONV_V3=1 timeout 60 python scripts/check_conv_v3.py > gpurun_out/check_v3.log 2>&1; echo "exit $?"; grep -v "^PASS" gpurun_out/check_v3.log | cut -c1-200
echo "== bench v3"; DCNN_CONV_V3=1 timeout 60 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_v3.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_v3.log | cut -c1-330
