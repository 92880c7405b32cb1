#!/bin/bash
# Offline subset build of the UNMODIFIED reference (SURVEY §6.2 plan (b)): the reference is a CMake C++/CUDA project,
# not a Python package, and its CMake needs network access (FetchContent: asio, json, zstd, gtest, stb) and TBB.
# This script copies /root/reference verbatim to baseline/_ref/src and compiles the non-pipeline libraries
# (device, ops, math, tensor, nn, accuracy) + baseline/ref_bench.cpp with nvcc, -DUSE_CUDA -DUSE_CUDNN, native sm_100
# SASS (the manual -DCUDA_ARCH override path of cmake/FindCUDA.cmake:83,96-99), cuDNN + cuBLAS + cuRAND.
# nlohmann/json.hpp comes from the copy bundled with the installed cudnn_frontend package.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF_SRC="${REF_SRC:-/root/reference}"
OUT="$HERE/_ref"
mkdir -p "$OUT/bin" "$OUT/obj"
if [ ! -d "$OUT/src" ]; then cp -r "$REF_SRC" "$OUT/src"; fi
S="$OUT/src"
JSON_INC="$(python - <<'PY'
import glob, sys, os, site
for d in site.getsitepackages():
    p = os.path.join(d, "include", "cudnn_frontend", "thirdparty")
    if os.path.exists(os.path.join(p, "nlohmann", "json.hpp")):
        print(p); break
PY
)"
CUDNN_INC=/usr/include/x86_64-linux-gnu
NVCC="${NVCC:-nvcc}"
FLAGS=(-O3 -std=c++17 -DUSE_CUDA -DUSE_CUDNN -DNDEBUG -gencode arch=compute_100,code=sm_100
       -I"$S/include" -I"$JSON_INC" -I"$CUDNN_INC" -Xcompiler -march=x86-64-v3,-fopenmp,-w -w --expt-relaxed-constexpr)
SRCS=$(find "$S/src/device" "$S/src/ops" "$S/src/math" "$S/src/tensor" "$S/src/nn" "$S/src/utils/accuracy_impl" "$S/src/cuda" \
       -name '*.cpp' -o -name '*.cu' | sort)
OBJS=()
pids=()
for f in $SRCS; do
  o="$OUT/obj/$(echo "${f#$S/}" | tr '/' '_').o"
  OBJS+=("$o")
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ]; then
    ( "$NVCC" "${FLAGS[@]}" -x cu -c "$f" -o "$o" ) &
    pids+=($!)
    if [ "${#pids[@]}" -ge 8 ]; then wait "${pids[0]}"; pids=("${pids[@]:1}"); fi
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
"$NVCC" "${FLAGS[@]}" -x cu -c "$HERE/ref_bench.cpp" -o "$OUT/obj/ref_bench.o"
"$NVCC" -o "$OUT/bin/ref_bench" "$OUT/obj/ref_bench.o" "${OBJS[@]}" -gencode arch=compute_100,code=sm_100 \
   -L/usr/lib/x86_64-linux-gnu -lcudnn -lcublas -lcurand -lgomp -Xlinker -rpath,/usr/lib/x86_64-linux-gnu
echo "built $OUT/bin/ref_bench"
