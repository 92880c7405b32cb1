#!/usr/bin/env python
"""GEMM benchmark (counterpart of benchmarks/gemm_benchmark.cpp: M=64, N=128, K=2 097 152 in NN/NT layouts) plus the
shapes ResNet-50 produces.  On CUDA it times this library's tcgen05 Dense kernels (dcnn_conv_tc as a 1x1 conv and
dcnn_wgrad_tc) with CUDA events and reports TFLOP/s against the measured peaks; on CPU it times ATen."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dcnn_b200 import ops  # noqa: E402


def timed(fn, iters, cuda):
    fn()
    if cuda:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    import time
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    return (time.perf_counter() - t0) * 1000 / iters


def main():
    cuda = torch.cuda.is_available()
    dev = "cuda" if cuda else "cpu"
    shapes = [(64, 128, 2097152 if cuda else 65536), (262144, 256, 64), (262144, 64, 256), (65536, 512, 128),
              (16384, 1024, 256), (4096, 2048, 512), (256, 200, 2048)]
    # the library GEMM of the same precision class next to ours (the reference's Dense layer is cuBLAS SGEMM,
    # src/math/cuda/gemm.cu:28): torch.matmul = cuBLAS / cuBLASLt with TF32 allowed, y = x w^T and dw = dy^T x
    torch.backends.cuda.matmul.allow_tf32 = True
    print(f"{'M':>8} {'N':>6} {'K':>8} | {'fwd ms':>9} {'TFLOP/s':>8} {'cuBLAS ms':>10} {'ratio':>6} | "
          f"{'wgrad ms':>9} {'TFLOP/s':>8} {'cuBLAS ms':>10} {'ratio':>6}")
    for M, N, K in shapes:
        x = torch.randn(M, K, device=dev)
        w = torch.randn(N, K, device=dev) / K ** 0.5
        dy = torch.randn(M, N, device=dev)
        dw = torch.zeros(N, K, device=dev)
        t_f = timed(lambda: ops.dense_forward(x, w, None), 10, cuda)
        t_w = timed(lambda: ops.dense_backward_weight(dy, x, dw), 10, cuda)
        c_f = timed(lambda: torch.matmul(x, w.t()), 10, cuda)
        c_w = timed(lambda: dw.addmm_(dy.t(), x), 10, cuda)
        fl = 2.0 * M * N * K
        print(f"{M:>8} {N:>6} {K:>8} | {t_f:>9.3f} {fl / t_f / 1e9:>8.1f} {c_f:>10.3f} {c_f / t_f:>5.2f}x | "
              f"{t_w:>9.3f} {fl / t_w / 1e9:>8.1f} {c_w:>10.3f} {c_w / t_w:>5.2f}x")
    print("ratio = library time / ours (> 1: ours is faster); back-to-back launches, operands larger than L2 except the "
          "last shape")


if __name__ == "__main__":
    main()
