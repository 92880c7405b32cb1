"""Stand-alone timing / ncu target for the tcgen05 convolution kernels on ResNet-50 (64x64, batch 256) layer shapes.
Normal run: CUDA-event timing, prints us and TFLOP/s per op.  NCU=1: launches each op once (3 x3 fprop, 1x1 fprop,
3x3 wgrad) so `ncu -k regex:conv_tc2|wgrad_tc -c 3` captures exactly one of each."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200.ops import cuda as K  # noqa: E402

CL = torch.channels_last
dev = "cuda:0"
ncu = os.environ.get("NCU") == "1"
torch.manual_seed(0)


def mk(n, c, h, w):
    return torch.randn(n, c, h, w, device=dev).contiguous(memory_format=CL)


def timed(name, fn, flops, iters=20, warm=5):
    if ncu:
        fn()
        torch.cuda.synchronize()
        return
    for _ in range(warm):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(iters):
        flush.zero_()                       # evict L2 (126 MB) between timed launches
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    med = ts[len(ts) // 2]
    print(f"{name:34s} median {med:8.1f} us  min {ts[0]:8.1f} us  {flops / med / 1e6:7.1f} TFLOP/s (tf32)", flush=True)


N = 256
# (name, Cin, Cout, H, R, pad)
shapes = [("3x3 128->128 @32x32", 128, 128, 32, 3, 1),
          ("1x1 256->1024 @16x16", 256, 1024, 16, 1, 0),
          ("3x3 256->256 @16x16", 256, 256, 16, 3, 1),
          ("1x1 512->128 @32x32", 512, 128, 32, 1, 0),
          ("3x3 512->512 @8x8", 512, 512, 8, 3, 1)]
if ncu:
    shapes = shapes[:2]
for name, ci, co, h, r, p in shapes:
    x = mk(N, ci, h, h)
    w = torch.randn(co, ci, r, r, device=dev).contiguous(memory_format=CL) * 0.05
    fl = 2.0 * N * h * h * co * ci * r * r
    timed("fprop+stats " + name, lambda: K.conv2d_forward(x, w, None, (1, 1), (p, p), want_stats=True), fl)
    if not ncu:
        dy = mk(N, co, h, h)
        timed("dgrad       " + name, lambda: K.conv2d_backward_data(dy, w, (N, ci, h, h), (1, 1), (p, p)), fl)
ci, co, h = 128, 128, 32
x, dy = mk(N, ci, h, h), mk(N, co, h, h)
dw = torch.zeros(co, ci, 3, 3, device=dev).contiguous(memory_format=CL)
timed("wgrad       3x3 128->128 @32x32", lambda: K.conv2d_backward_weight(dy, x, dw, (1, 1), (1, 1)),
      2.0 * N * h * h * co * ci * 9)
if not ncu:
    ci, co, h = 256, 1024, 16
    x, dy = mk(N, ci, h, h), mk(N, co, h, h)
    dw = torch.zeros(co, ci, 1, 1, device=dev).contiguous(memory_format=CL)
    timed("wgrad       1x1 256->1024 @16x16", lambda: K.conv2d_backward_weight(dy, x, dw, (1, 1), (0, 0)),
          2.0 * N * h * h * co * ci)
    # copy bandwidth reference for the roofline denominator
    a = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    b = torch.empty_like(a)
    for _ in range(3):
        b.copy_(a)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        b.copy_(a)
    e1.record()
    e1.synchronize()
    print(f"d2d copy 1 GiB: {2 * 10 * (1 << 30) / (e0.elapsed_time(e1) * 1e-3) / 1e12:.2f} TB/s (read+write)")
print("done")
