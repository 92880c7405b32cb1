#!/usr/bin/env python
"""Per-op benchmark (counterpart of benchmarks/tensor_ops_benchmark.cpp, which times im2col/col2im and NCHW<->CNHW on
CPU vs CUDA).  Those ops do not exist here (implicit GEMM through TMA), so this times what replaced them: conv
fprop/dgrad/wgrad per ResNet-50 layer shape, and the BatchNorm / pooling passes, with achieved TFLOP/s or GB/s."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dcnn_b200 import ops  # noqa: E402
from gemm_benchmark import timed  # noqa: E402


def main():
    cuda = torch.cuda.is_available()
    dev = "cuda" if cuda else "cpu"
    B = int(os.environ.get("BATCH", 256 if cuda else 8))
    convs = [(64, 32, 64, 1, 1), (64, 32, 64, 3, 1), (64, 32, 256, 1, 1), (256, 32, 128, 1, 1), (128, 32, 128, 3, 2),
             (128, 16, 512, 1, 1), (256, 8, 256, 3, 1), (1024, 8, 256, 1, 1), (512, 4, 512, 3, 1), (512, 4, 2048, 1, 1)]
    print(f"batch {B}\n{'conv (C,HW,OC,k,s)':<24} {'fprop ms':>9} {'dgrad ms':>9} {'wgrad ms':>9} {'fprop TFLOP/s':>14}")
    for C, HW, OC, k, s in convs:
        p = k // 2
        x = torch.randn(B, C, HW, HW, device=dev).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(OC, C, k, k, device=dev) / (C * k * k) ** 0.5).contiguous(memory_format=torch.channels_last)
        y, _ = ops.conv2d_forward(x, w, None, (s, s), (p, p))
        dy = torch.randn_like(y)
        dw = torch.zeros_like(w)
        tf = timed(lambda: ops.conv2d_forward(x, w, None, (s, s), (p, p), True), 10, cuda)
        td = timed(lambda: ops.conv2d_backward_data(dy, w, tuple(x.shape), (s, s), (p, p)), 10, cuda)
        tw = timed(lambda: ops.conv2d_backward_weight(dy, x, dw, (s, s), (p, p)), 10, cuda)
        fl = 2.0 * y.numel() * C * k * k
        print(f"{str((C, HW, OC, k, s)):<24} {tf:>9.3f} {td:>9.3f} {tw:>9.3f} {fl / tf / 1e9:>14.1f}")
    print(f"\n{'BN (C,HW)':<24} {'fwd ms':>9} {'GB/s':>8} {'bwd ms':>9} {'GB/s':>8}")
    for C, HW in [(64, 32), (256, 32), (512, 16), (1024, 8), (2048, 4)]:
        x = torch.randn(B, C, HW, HW, device=dev).contiguous(memory_format=torch.channels_last)
        g, b = torch.ones(C, 1, 1, 1, device=dev), torch.zeros(C, 1, 1, 1, device=dev)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        dg, db = torch.zeros_like(g), torch.zeros_like(b)

        def fwd():
            m, i = ops.bn_finalize(ops.bn_stats(x), 1e-5, 0.1, rm, rv, gamma=g, beta=b)
            return m, i, ops.bn_apply(x, m, i, g, b, None, True)
        m, i, out = fwd()
        dout = torch.randn_like(x)
        t1 = timed(fwd, 10, cuda)
        t2 = timed(lambda: ops.bn_backward(dout, x, out, m, i, g, dg, db, True, False), 10, cuda)
        nbytes = x.numel() * 4
        print(f"{str((C, HW)):<24} {t1:>9.3f} {3 * nbytes / t1 / 1e6:>8.0f} {t2:>9.3f} {7 * nbytes / t2 / 1e6:>8.0f}")


if __name__ == "__main__":
    main()
