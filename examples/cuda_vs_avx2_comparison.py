#!/usr/bin/env python
"""Counterpart of the reference's examples/cuda_vs_avx2_comparison.cpp: the same GEMM on this library's AVX2 CPU kernels,
on ATen's CPU GEMM and (when a GPU is present) on the tcgen05 Dense kernel, with the results cross-checked."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dcnn_b200.ops import cpu_native, tensor_ops  # noqa: E402


def main():
    M, N, K = 1024, 512, 1024
    a, b = torch.randn(M, K), torch.randn(K, N)
    t0 = time.perf_counter()
    c_native = cpu_native.gemm(a, b)
    t_native = time.perf_counter() - t0
    t0 = time.perf_counter()
    c_aten = a @ b
    t_aten = time.perf_counter() - t0
    err = float((c_native - c_aten).abs().max() / c_aten.abs().max())
    fl = 2.0 * M * N * K / 1e9
    print(f"sgemm {M}x{N}x{K}: native AVX2 {t_native * 1e3:.2f} ms ({fl / t_native:.0f} GFLOP/s), "
          f"ATen {t_aten * 1e3:.2f} ms ({fl / t_aten:.0f} GFLOP/s), max rel diff {err:.2e}")
    if torch.cuda.is_available():
        ag, bg = a.cuda(), b.cuda()
        tensor_ops.gemm(ag, bg)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        cg = tensor_ops.gemm(ag, bg)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        errg = float((cg.cpu() - c_aten).abs().max() / c_aten.abs().max())
        print(f"tcgen05 (TF32) on the GPU: {ms:.3f} ms ({fl / ms:.1f} TFLOP/s), max rel diff vs FP32 {errg:.2e}")
    else:
        print("no CUDA device: GPU arm skipped")


if __name__ == "__main__":
    main()
