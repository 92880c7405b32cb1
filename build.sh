#!/usr/bin/env bash
# Build front end with the reference's switches (its build.sh: --clean, --debug, --verbose, --cuda, --mkl).
#   ./build.sh              device library (nvcc, sm_100a) + host runtime (g++), in-tree under dcnn_b200/lib
#   ./build.sh --clean      remove build products first
#   ./build.sh --host-only  host runtime only (no nvcc needed)
#   ./build.sh --debug      additionally build and run the sanitizer self-tests of the host runtime (ASan/UBSan, TSan)
#   ./build.sh --verbose    show ptxas resource usage per kernel
#   ./build.sh --reference  also build the offline subset of the reference used as benchmark arm and test oracle
# --cuda is accepted and implied (there is no non-CUDA device build); --mkl is accepted and ignored (the CPU path uses
# this library's own AVX2 / AVX-512 kernels or ATen).
set -euo pipefail
root="$(cd "$(dirname "$0")" && pwd)"
cd "$root"
clean=0 debug=0 verbose=0 host_only=0 reference=0
for a in "$@"; do
  case "$a" in
    --clean) clean=1 ;; --debug) debug=1 ;; --verbose) verbose=1 ;; --host-only) host_only=1 ;;
    --reference) reference=1 ;; --cuda|--mkl) ;;
    -h|--help) sed -n 2,11p "$0"; exit 0 ;;
    *) echo "unknown option $a" >&2; exit 2 ;;
  esac
done
if [ "$clean" = 1 ]; then
  rm -rf build build_sanitize dcnn_b200/lib/*.so dcnn_b200/lib/*.sha256
  echo "cleaned"
fi
python - <<PY
from dcnn_b200.build import build, build_host
print("host runtime:", build_host(force=bool($clean)))
if not $host_only:
    print("device library:", build(force=bool($clean), verbose=bool($verbose)))
PY
if [ "$debug" = 1 ]; then bash scripts/sanitize_host.sh "$root/build_sanitize"; fi
if [ "$reference" = 1 ]; then bash baseline/build_ref.sh; fi
