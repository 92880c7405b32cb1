mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
echo "== dp tests"; timeout 900 python -m pytest tests/test_gpu_dp.py -m gpu -x -q --timeout 400 -p no:cacheprovider > gpurun_out/t_dp.log 2>&1; echo "exit $?"; tail -15 gpurun_out/t_dp.log | cut -c1-400


