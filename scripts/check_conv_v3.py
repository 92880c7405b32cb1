"""Numerics + timing check of the 256-pixel-tile convolution kernel (conv_tc3; DCNN_CONV_V3=1) or, with ``--pair`` and
DCNN_CONV_2CTA=1, of the CTA-pair kernel (conv_tc5).  Run in a fresh process: the switches are read once.  Compares fprop (+ epilogue BN statistics), dgrad and dgrad-accumulate against
cuDNN FP32 (allow_tf32 off) on shapes large enough to select the kernel, including partial tiles; prints PASS/FAIL and
CUDA-event timings with an L2 flush between launches.  Exit code 0 only if everything passed."""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200._lib import I, lib  # noqa: E402
from dcnn_b200.ops import cuda as K  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
CL = torch.channels_last
dev = "cuda:0"
torch.manual_seed(0)
ok = True


def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


def check(name, e, tol):
    global ok
    good = e < tol
    ok &= good
    print(f"{'PASS' if good else 'FAIL'} {name}: rel-L2 {e:.2e} (tol {tol:.0e})", flush=True)


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
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


# (N, Cin, H, W, Cout, k, stride, pad, bias)
PAIR_CASES = [(64, 64, 32, 32, 256, 1, 1, 0, True),      # DCNN_CONV_2CTA=1: CTA-pair kernel, one 256-wide N tile
              (37, 32, 15, 15, 320, 3, 1, 1, False),     # partial pixel tiles, partial second N tile (320 = 256 + 64)
              (149, 64, 8, 8, 512, 3, 1, 1, True),       # odd number of 128-pixel tiles: the last pair has an empty half
              (32, 256, 16, 16, 1024, 1, 1, 0, False)]   # ResNet-50 stage-2 expand shape, four N tiles
cases = PAIR_CASES if "--pair" in sys.argv else [(64, 64, 32, 32, 128, 3, 1, 1, True),     # 256 full tiles, two N tiles? no: Cout 128 -> one
         (40, 32, 33, 31, 64, 3, 1, 1, False),      # partial tiles in h and w, BN=64 variant
         (48, 64, 32, 32, 320, 1, 1, 0, True),      # three N tiles, the last one partial (320 = 2.5 x 128)
         (602, 32, 16, 16, 128, 3, 2, 1, False),    # strided: element-stride box, 8x8 outputs, box spans 4 images (last partial)
         (64, 128, 32, 32, 128, 3, 1, 1, False)]
for (N, Ci, H, W, Co, k, s, p, bias) in cases:
    oh, ow = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
    tiles = lib().dcnn_conv_tc_stat_tiles(I(N), I(oh), I(ow), I(s), I(s), I(Co), I(k * k * ((Ci + 31) // 32)))
    base = lib().dcnn_conv_tc_m_tiles(I(N), I(oh), I(ow), I(s), I(s))
    tag = f"[{N}x{Ci}x{H}x{W} -> {Co} k{k} s{s}] stat tiles {tiles} (v2 would use {base})"
    print(tag, flush=True)
    x = torch.randn(N, Ci, H, W, device=dev).contiguous(memory_format=CL)
    w = (torch.randn(Co, Ci, k, k, device=dev) / (Ci * k * k) ** 0.5).contiguous(memory_format=CL)
    b = torch.randn(Co, device=dev) if bias else None
    y_ref = F.conv2d(x, w, b, stride=s, padding=p)
    y, st = K.conv2d_forward(x, w, None if b is None else b.reshape(Co, 1, 1, 1), (s, s), (p, p), want_stats=True)
    check("fprop", rel(y, y_ref), 2e-3)
    rm, rv = torch.zeros(Co, device=dev), torch.ones(Co, device=dev)
    mean, inv = K.bn_finalize(st, 1e-5, 0.1, rm, rv)
    check("epilogue mean", float((mean - y_ref.mean((0, 2, 3))).abs().max() / (y_ref.std() + 1e-12)), 2e-3)
    inv_ref = (y_ref.var((0, 2, 3), unbiased=False) + 1e-5).rsqrt()
    check("epilogue invstd", rel(inv[0], inv_ref), 5e-3)
    if s == 1:
        dy = torch.randn(N, Co, oh, ow, device=dev).contiguous(memory_format=CL)
        dx_ref = torch.nn.grad.conv2d_input((N, Ci, H, W), w, dy, stride=s, padding=p)
        dx = K.conv2d_backward_data(dy, w, (N, Ci, H, W), (s, s), (p, p))
        check("dgrad", rel(dx, dx_ref), 2e-3)
        acc = torch.randn(N, Ci, H, W, device=dev).contiguous(memory_format=CL)
        want = acc + dx_ref
        K.conv2d_backward_data(dy, w, (N, Ci, H, W), (s, s), (p, p), out=acc, accumulate=True)
        check("dgrad accumulate", rel(acc, want), 2e-3)
    torch.cuda.synchronize()

N = 256
TIMED = [("3x3 128->128 @32x32", 128, 128, 32, 3, 1), ("1x1 256->1024 @16x16", 256, 1024, 16, 1, 0),
         ("3x3 256->256 @16x16", 256, 256, 16, 3, 1), ("1x1 512->128 @32x32", 512, 128, 32, 1, 0)]
if "--pair" in sys.argv:
    TIMED = [("1x1 256->1024 @16x16", 256, 1024, 16, 1, 0), ("3x3 256->256 @16x16", 256, 256, 16, 3, 1),
             ("1x1 128->512 @16x16", 128, 512, 16, 1, 0), ("1x1 64->256 @32x32", 64, 256, 32, 1, 0),
             ("1x1 512->2048 @4x4", 512, 2048, 4, 1, 0)]
for name, ci, co, h, r, p in TIMED:
    x = torch.randn(N, ci, h, h, device=dev).contiguous(memory_format=CL)
    w = (torch.randn(co, ci, r, r, device=dev) * 0.05).contiguous(memory_format=CL)
    us = timed(lambda: K.conv2d_forward(x, w, None, (1, 1), (p, p), want_stats=True))
    print(f"time fprop+stats {name:24s} {us:8.1f} us  {2.0 * N * h * h * co * ci * r * r / us / 1e6:7.1f} TFLOP/s",
          flush=True)
print("ALL PASS" if ok else "SOME FAILED")
sys.exit(0 if ok else 1)
