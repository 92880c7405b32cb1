mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() { # name nproc args...
  name=$1; n=$2; shift 2
  echo "== $name"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $n --steps 10 --warmup 4 "$@" > gpurun_out/bench_$name.log 2>&1; echo "exit $?"
  grep '^{' gpurun_out/bench_$name.log | python -c "import sys,json; [print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['config']['parallelism'], 'bubble', d.get('pipeline_bubble_fraction'), 'e2e', d['e2e']['value'], 'loss', d['final_loss']) for d in map(json.loads, sys.stdin)]" || tail -5 gpurun_out/bench_$name.log
}
run dp4 4 --parallelism dp4
run pp4 4 --parallelism pp4
SCHEDULE=sync run pp4_sync 4 --parallelism pp4
run pp2xdp2 4 --parallelism pp2xdp2
run pp4_mb16 4 --parallelism pp4 --microbatches 16
