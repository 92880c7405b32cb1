mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
echo "== conv+dense+bn tests"; timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 120 -k "conv or dense or batchnorm" -p no:cacheprovider > gpurun_out/t_conv2.log 2>&1; echo "exit $?"; tail -12 gpurun_out/t_conv2.log | cut -c1-300
echo "== models"; timeout 600 python -m pytest tests/test_gpu_models.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/t_models2.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_models2.log | cut -c1-300
echo "== bench graph"; timeout 400 python bench.py --steps 10 --warmup 4 > gpurun_out/bench_graph2.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_graph2.log | cut -c1-400
echo "== bench v1"; DCNN_CONV_V1=1 timeout 400 python bench.py --steps 10 --warmup 4 > gpurun_out/bench_graph_v1.log 2>&1; echo "exit $?"; tail -1 gpurun_out/bench_graph_v1.log | cut -c1-250
