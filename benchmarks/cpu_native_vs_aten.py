#!/usr/bin/env python
"""Native AVX2 GEMM / elementwise kernels (csrc/host/cpu_math.cpp) against ATen on this host -- the counterpart of the
reference's benchmarks/cuda_vs_avx2_comparison.cpp and the CPU half of gemm_benchmark.cpp.  Wall-clock, best of N."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dcnn_b200.ops import cpu_native as N  # noqa: E402


def best(fn, reps=5):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts)


def main():
    threads = int(os.environ.get("DCNN_CPU_THREADS", torch.get_num_threads()))
    torch.set_num_threads(threads)
    N.set_num_threads(threads)
    print(f"threads: {threads}")
    print(f"{'op':34s} {'native ms':>10s} {'ATen ms':>10s} {'native GFLOP/s':>15s} {'ATen GFLOP/s':>13s}")
    for M, Nn, K in [(256, 256, 256), (512, 512, 512), (1024, 1024, 1024), (64, 128, 65536), (4096, 64, 576),
                     (128, 10, 1024)]:
        a, b = torch.randn(M, K), torch.randn(K, Nn)
        c = torch.empty(M, Nn)
        tn = best(lambda: N.gemm(a, b, c=c))
        ta = best(lambda: torch.mm(a, b, out=c))
        fl = 2.0 * M * Nn * K / 1e9
        print(f"{'sgemm %dx%dx%d' % (M, Nn, K):34s} {tn * 1e3:10.3f} {ta * 1e3:10.3f} {fl / tn:15.1f} {fl / ta:13.1f}")
    n = 1 << 24
    x, y = torch.randn(n), torch.randn(n)
    out = torch.empty(n)
    for name, fn_n, fn_a in [("add 16M", lambda: N.binary("add", x, y, out), lambda: torch.add(x, y, out=out)),
                             ("relu 16M", lambda: N.unary("relu", x, out), lambda: torch.relu(x)),
                             ("dot 16M", lambda: N.dot(x, y), lambda: torch.dot(x, y))]:
        tn, ta = best(fn_n), best(fn_a)
        print(f"{name:34s} {tn * 1e3:10.3f} {ta * 1e3:10.3f}")


if __name__ == "__main__":
    main()
