#!/usr/bin/env python
"""This library's tcgen05 convolution kernels against cuDNN (through torch, channels_last, TF32 allowed -- the same
precision class) on the ResNet-50 64x64 batch-256 layer shapes: fprop (+ BN statistics on our side, which cuDNN does
not produce), dgrad, wgrad.  CUDA events, L2 flushed between launches, median of 20.  (the bar is
'>= cuDNN on the same box'.)"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200.ops import cuda as K  # noqa: E402

torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True
CL = torch.channels_last
dev = "cuda:0"
flush = None


def timed(fn, iters=20, warm=5):
    global flush
    if flush is None:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    N = int(os.environ.get("BENCH_BATCH", 256))
    # (Cin, Cout, H, k, stride, pad): the distinct conv shapes of ResNet-50 on 64x64 inputs after the stem + max-pool
    shapes = [(64, 64, 32, 1, 1, 0), (64, 64, 32, 3, 1, 1), (64, 256, 32, 1, 1, 0), (256, 64, 32, 1, 1, 0),
              (256, 128, 32, 1, 1, 0), (128, 128, 32, 3, 2, 1), (128, 512, 16, 1, 1, 0), (512, 128, 16, 1, 1, 0),
              (128, 128, 16, 3, 1, 1), (512, 256, 16, 1, 1, 0), (256, 256, 16, 3, 2, 1), (256, 1024, 8, 1, 1, 0),
              (1024, 256, 8, 1, 1, 0), (256, 256, 8, 3, 1, 1), (1024, 512, 8, 1, 1, 0), (512, 512, 8, 3, 2, 1),
              (512, 2048, 4, 1, 1, 0), (2048, 512, 4, 1, 1, 0), (512, 512, 4, 3, 1, 1)]
    if os.environ.get("BENCH_SHAPES"):      # comma-separated indices into the list above
        shapes = [shapes[int(i)] for i in os.environ["BENCH_SHAPES"].split(",")]
    print(f"batch {N}; times in us (ours / cuDNN), ratio > 1 means ours is faster")
    print(f"{'shape':32s} {'fprop':>17s} {'dgrad':>17s} {'wgrad':>17s}")
    tot = [0.0, 0.0]
    for ci, co, h, k, s, p in shapes:
        oh = (h + 2 * p - k) // s + 1
        x = torch.randn(N, ci, h, h, device=dev).contiguous(memory_format=CL)
        w = (torch.randn(co, ci, k, k, device=dev) * 0.05).contiguous(memory_format=CL)
        dy = torch.randn(N, co, oh, oh, device=dev).contiguous(memory_format=CL)
        dw = torch.zeros_like(w)
        cells = []
        for ours, ref in (
            (lambda: K.conv2d_forward(x, w, None, (s, s), (p, p), want_stats=True),
             lambda: F.conv2d(x, w, None, stride=s, padding=p)),
            (lambda: K.conv2d_backward_data(dy, w, (N, ci, h, h), (s, s), (p, p)),
             lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (s, s), (p, p), (1, 1), False, (0, 0), 1,
                                                         (True, False, False))),
            (lambda: K.conv2d_backward_weight(dy, x, dw, (s, s), (p, p)),
             lambda: torch.ops.aten.convolution_backward(dy, x, w, None, (s, s), (p, p), (1, 1), False, (0, 0), 1,
                                                         (False, True, False))),
        ):
            a, b = timed(ours), timed(ref)
            tot[0] += a
            tot[1] += b
            cells.append(f"{a:6.1f}/{b:6.1f} {b / a:4.2f}x")
        print(f"{'%dx%d %d->%d @%d s%d' % (k, k, ci, co, h, s):32s} " + " ".join(f"{c:>17s}" for c in cells), flush=True)
    print(f"sum over shapes: ours {tot[0] / 1e3:.2f} ms, cuDNN {tot[1] / 1e3:.2f} ms, ratio {tot[1] / tot[0]:.2f}x")


if __name__ == "__main__":
    main()
