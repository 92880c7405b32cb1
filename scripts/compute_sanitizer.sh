#!/usr/bin/env bash
# compute-sanitizer over the GPU kernel tests (run on a GPU box, ONE tool per invocation -- each replays every kernel):
#   scripts/compute_sanitizer.sh memcheck  [pytest -k expression]
#   scripts/compute_sanitizer.sh racecheck [pytest -k expression]      (shared-memory hazards)
#   scripts/compute_sanitizer.sh synccheck [pytest -k expression]      (barrier misuse)
#   scripts/compute_sanitizer.sh initcheck [pytest -k expression]      (reads of uninitialised device memory)
# The flag-polling multi-GPU kernels (csrc/comm.cu, *_sync in csrc/norm.cu) are excluded: under the sanitizer's
# serialisation their peers do not make progress.
set -euo pipefail
tool="${1:-memcheck}"
expr="${2:-conv or dense or batchnorm or pool or loss or optim}"
mkdir -p gpurun_out
exec timeout 1500 compute-sanitizer --tool "$tool" --error-exitcode 3 --launch-timeout 120 \
    --log-file "gpurun_out/sanitizer_${tool}.log" \
    python -m pytest tests/test_gpu_ops.py -m gpu -x -q -p no:cacheprovider -k "$expr"
