mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
echo "== dp tests"; timeout 900 python -m pytest tests/test_gpu_dp.py -m gpu -x -q --timeout 400 -p no:cacheprovider > gpurun_out/t_dp.log 2>&1; echo "exit $?"; tail -15 gpurun_out/t_dp.log | cut -c1-400
echo "== bench dp2 eager"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 4 --no-graph > gpurun_out/bench_dp2_eager.log 2>&1; echo "exit $?"; tail -2 gpurun_out/bench_dp2_eager.log | cut -c1-500
echo "== bench dp2 graph"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 4 > gpurun_out/bench_dp2_graph.log 2>&1; echo "exit $?"; tail -2 gpurun_out/bench_dp2_graph.log | cut -c1-500
