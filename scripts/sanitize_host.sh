#!/usr/bin/env bash
# AddressSanitizer + UndefinedBehaviorSanitizer run of the native host runtime (csrc/host/*.cpp) through its self-test
# driver.  Usage: scripts/sanitize_host.sh [build-dir]   (default: ./build_sanitize)
set -euo pipefail
root="$(cd "$(dirname "$0")/.." && pwd)"
out="${1:-$root/build_sanitize}"
mkdir -p "$out"
g++ -std=c++17 -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined \
    -fopenmp -march=x86-64-v3 -I "$root/csrc/host" \
    "$root"/csrc/host/*.cpp "$root/csrc/host_selftest/selftest_main.cpp" "$root/csrc/host_selftest/selftest_tcp.cpp" \
    -o "$out/host_selftest"
ASAN_OPTIONS=detect_leaks=1:abort_on_error=0 UBSAN_OPTIONS=print_stacktrace=1 OMP_NUM_THREADS=4 "$out/host_selftest" "$out"
# ThreadSanitizer run of the TCP transport alone (io threads, sender threads, consumer threads, teardown)
g++ -std=c++17 -O1 -g -fno-omit-frame-pointer -fsanitize=thread -DSELFTEST_TCP_MAIN -I "$root/csrc/host" \
    "$root/csrc/host/tcp_transport.cpp" "$root/csrc/host_selftest/selftest_tcp.cpp" -o "$out/tcp_selftest_tsan" -lpthread
TSAN_OPTIONS=halt_on_error=0:second_deadlock_stack=1 "$out/tcp_selftest_tsan"
