#!/bin/bash
# Offline subset build of the UNMODIFIED reference (SURVEY §6.2 plan (b)): the reference is a CMake C++/CUDA project,
# not a Python package, and its CMake needs network access (FetchContent: asio, json, zstd, gtest, stb) and TBB.
# This script copies /root/reference verbatim to baseline/_ref/src and compiles the non-pipeline libraries
# (device, ops, math, tensor, nn, accuracy) + baseline/ref_bench.cpp: g++ -O3 -march=x86-64-v3 for .cpp (the reference's
# own release flags minus LTO), nvcc for .cu with native sm_100 SASS (the manual -DCUDA_ARCH override path of
# cmake/FindCUDA.cmake:83,96-99), -DUSE_CUDA -DUSE_CUDNN, linked against cuDNN + cuBLAS + cuRAND.
# Threading: no TBB/OpenMP (serial parallel_for, thread_handler.hpp:88-90) - irrelevant for the GPU path.
# nlohmann/json.hpp comes from the copy bundled with the installed cudnn_frontend package.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF_SRC="${REF_SRC:-/root/reference}"
OUT="$HERE/_ref"
mkdir -p "$OUT/bin" "$OUT/obj"
if [ ! -d "$OUT/src" ]; then cp -r "$REF_SRC" "$OUT/src"; fi
S="$OUT/src"
JSON_INC="$(python - <<'PY'
import os, site
for d in site.getsitepackages():
    p = os.path.join(d, "include", "cudnn_frontend", "thirdparty")
    if os.path.exists(os.path.join(p, "nlohmann", "json.hpp")):
        print(p); break
PY
)"
CUDA_HOME="${CUDA_HOME:-/usr/local/cuda}"
CUDNN_INC=/usr/include/x86_64-linux-gnu
NVCC="${NVCC:-$CUDA_HOME/bin/nvcc}"
DEFS=(-DUSE_CUDA -DUSE_CUDNN -DNDEBUG)
INCS=(-I"$S/include" -I"$JSON_INC" -I"$CUDNN_INC" -I"$CUDA_HOME/include")
# -include immintrin.h: src/math/cpu/{s,d}gemm.cpp use AVX2 intrinsics without including the header (they get it
# transitively from TBB headers in the stock build); force-including it keeps the sources byte-identical.
CXXFLAGS=(-O3 -std=c++17 -march=x86-64-v3 -w -fpermissive -include immintrin.h -include iostream "${DEFS[@]}" "${INCS[@]}")
NVFLAGS=(-O3 -std=c++17 -gencode arch=compute_100,code=sm_100 -w --expt-relaxed-constexpr -include iostream "${DEFS[@]}" "${INCS[@]}"
         -Xcompiler -march=x86-64-v3)
SRCS=$(find "$S/src/device" "$S/src/ops" "$S/src/math" "$S/src/tensor" "$S/src/nn" "$S/src/utils/accuracy_impl" "$S/src/cuda" \
       \( -name '*.cpp' -o -name '*.cu' \) | sort)
OBJS=()
running=0
for f in $SRCS; do
  o="$OUT/obj/$(echo "${f#$S/}" | tr '/' '_').o"
  OBJS+=("$o")
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ]; then
    if [[ "$f" == *.cu ]]; then ( "$NVCC" "${NVFLAGS[@]}" -c "$f" -o "$o" ) &
    else ( g++ "${CXXFLAGS[@]}" -c "$f" -o "$o" ) & fi
    running=$((running+1))
    if [ "$running" -ge 8 ]; then wait -n; running=$((running-1)); fi
  fi
done
wait
for o in "${OBJS[@]}"; do [ -f "$o" ] || { echo "missing object $o"; exit 1; }; done
g++ "${CXXFLAGS[@]}" -c "$HERE/ref_bench.cpp" -o "$OUT/obj/ref_bench.o"
# shared cudart like the CMake build (CUDA::cudart): the DeviceManager singleton creates its stream during static
# initialisation, which is only safe when the runtime library is already loaded
"$NVCC" --cudart shared -o "$OUT/bin/ref_bench" "$OUT/obj/ref_bench.o" "${OBJS[@]}" -gencode arch=compute_100,code=sm_100 \
   -L/usr/lib/x86_64-linux-gnu -lcudnn -lcublas -lcurand -Xlinker -rpath,/usr/lib/x86_64-linux-gnu -Xlinker -rpath,/usr/local/cuda/lib64
# BASELINE config 1 (MNIST CNN on the CPU): same objects, second harness
g++ "${CXXFLAGS[@]}" -c "$HERE/ref_mnist_cpu.cpp" -o "$OUT/obj/ref_mnist_cpu.o"
OBJS_NOBENCH=()
for o in "${OBJS[@]}"; do OBJS_NOBENCH+=("$o"); done
"$NVCC" --cudart shared -o "$OUT/bin/ref_mnist_cpu" "$OUT/obj/ref_mnist_cpu.o" "${OBJS_NOBENCH[@]}" -gencode arch=compute_100,code=sm_100 \
   -L/usr/lib/x86_64-linux-gnu -lcudnn -lcublas -lcurand -Xlinker -rpath,/usr/lib/x86_64-linux-gnu -Xlinker -rpath,/usr/local/cuda/lib64
# snapshot interoperability harness (tests/test_reference_interop.py)
g++ "${CXXFLAGS[@]}" -c "$HERE/ref_snapshot.cpp" -o "$OUT/obj/ref_snapshot.o"
"$NVCC" --cudart shared -o "$OUT/bin/ref_snapshot" "$OUT/obj/ref_snapshot.o" "${OBJS_NOBENCH[@]}" -gencode arch=compute_100,code=sm_100 \
   -L/usr/lib/x86_64-linux-gnu -lcudnn -lcublas -lcurand -Xlinker -rpath,/usr/lib/x86_64-linux-gnu -Xlinker -rpath,/usr/local/cuda/lib64
# wire-format harness: the reference's BinarySerializer / Message / TBuffer (header-only, no asio or TBB needed)
g++ "${CXXFLAGS[@]}" -c "$HERE/ref_wire.cpp" -o "$OUT/obj/ref_wire.o"
"$NVCC" --cudart shared -o "$OUT/bin/ref_wire" "$OUT/obj/ref_wire.o" "${OBJS_NOBENCH[@]}" -gencode arch=compute_100,code=sm_100 \
   -L/usr/lib/x86_64-linux-gnu -lcudnn -lcublas -lcurand -Xlinker -rpath,/usr/lib/x86_64-linux-gnu -Xlinker -rpath,/usr/local/cuda/lib64
# data-path harness: the reference's loaders and NormalizationAugmentation
g++ "${CXXFLAGS[@]}" -c "$HERE/ref_data.cpp" -o "$OUT/obj/ref_data.o"
"$NVCC" --cudart shared -o "$OUT/bin/ref_data" "$OUT/obj/ref_data.o" "${OBJS_NOBENCH[@]}" -gencode arch=compute_100,code=sm_100 \
   -L/usr/lib/x86_64-linux-gnu -lcudnn -lcublas -lcurand -Xlinker -rpath,/usr/lib/x86_64-linux-gnu -Xlinker -rpath,/usr/local/cuda/lib64
echo "built $OUT/bin/ref_bench, $OUT/bin/ref_mnist_cpu, $OUT/bin/ref_snapshot, $OUT/bin/ref_wire and $OUT/bin/ref_data"
