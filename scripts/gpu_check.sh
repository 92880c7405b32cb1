mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
export PYTHONUNBUFFERED=1
echo "== probes"; timeout 300 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 120 -k "test_tma" -p no:cacheprovider > gpurun_out/t_probe.log 2>&1; echo "exit $?"; tail -5 gpurun_out/t_probe.log
echo "== elementwise"; timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 120 -k "not tma and not conv and not dense" -p no:cacheprovider > gpurun_out/t_elem.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_elem.log
echo "== conv fwd"; timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 120 -k "conv_forward or fp32_exact" -p no:cacheprovider > gpurun_out/t_convf.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_convf.log
echo "== conv dgrad"; timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 120 -k "conv_backward_data" -p no:cacheprovider > gpurun_out/t_convd.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_convd.log
echo "== conv wgrad"; timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 120 -k "conv_backward_weight" -p no:cacheprovider > gpurun_out/t_convw.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_convw.log
echo "== dense"; timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q --timeout 120 -k "dense" -p no:cacheprovider > gpurun_out/t_dense.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_dense.log
echo "== models (SIMT only)"; DCNN_DISABLE_TC=1 timeout 600 python -m pytest tests/test_gpu_models.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/t_models_simt.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_models_simt.log
echo "== models (TC)"; timeout 600 python -m pytest tests/test_gpu_models.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/t_models_tc.log 2>&1; echo "exit $?"; tail -8 gpurun_out/t_models_tc.log
echo "== bench eager"; timeout 400 python bench.py --steps 5 --warmup 3 --no-graph > gpurun_out/bench_eager.log 2>&1; echo "exit $?"; tail -3 gpurun_out/bench_eager.log
echo "== bench graph"; timeout 400 python bench.py --steps 10 --warmup 4 > gpurun_out/bench_graph.log 2>&1; echo "exit $?"; tail -3 gpurun_out/bench_graph.log
