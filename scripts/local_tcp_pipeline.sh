#!/usr/bin/env bash
# Start N `network_worker` processes on localhost ports and a coordinator that names them host:port -- the reference's
# way of running a pipeline without a cluster (N network_worker processes on ports 8001.. + a coordinator).
#   scripts/local_tcp_pipeline.sh [coordinator example] [number of stages] [first port] [worker flags...]
#   scripts/local_tcp_pipeline.sh examples/semi_async_pipeline_coordinator.py 2 8001 --gpu
set -euo pipefail
root="$(cd "$(dirname "$0")/.." && pwd)"
coord="${1:-$root/examples/sync_pipeline_coordinator.py}"
stages="${2:-2}"
port0="${3:-8001}"
shift $(( $# < 3 ? $# : 3 ))
pids=()
workers=""
cleanup() { for p in "${pids[@]}"; do kill "$p" 2>/dev/null || true; done; }
trap cleanup EXIT
for i in $(seq 0 $((stages - 1))); do
  port=$((port0 + i))
  if [[ " $* " == *" --gpu "* ]]; then dev=(--device "$i"); else dev=(); fi
  python "$root/examples/network_worker.py" "$port" --host 127.0.0.1 "${dev[@]}" "$@" &
  pids+=($!)
  workers="${workers:+$workers,}127.0.0.1:$port"
done
python "$coord" --workers "$workers"
wait "${pids[@]}"
pids=()
