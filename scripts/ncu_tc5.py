"""ncu target for the CTA-pair convolution kernel: one forced-variant launch each of conv_tc5 (cta_group::2, 256x256 pair
tiles) and conv_tc3 (256x128 tiles) on the same layer, plus CUDA-event timings of both when run plain."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200._lib import check, lib  # noqa: E402
from dcnn_b200.ops import cuda as K  # noqa: E402

CL = torch.channels_last
N, ci, co, h, r, p = 256, 256, 256, 16, 3, 1
shape = os.environ.get("TC5_SHAPE")
if shape:
    N, ci, co, h, r, p = map(int, shape.split(","))
x = torch.randn(N, ci, h, h, device="cuda").contiguous(memory_format=CL)
w = (torch.randn(co, ci, r, r, device="cuda") * 0.05).contiguous(memory_format=CL)


def run(variant):
    check(lib().dcnn_conv_tc_set_variant(C.c_int(variant)), "variant")
    try:
        return K.conv2d_forward(x, w, None, (1, 1), (p, p), want_stats=True)
    finally:
        check(lib().dcnn_conv_tc_set_variant(C.c_int(0)), "variant")


for v in (3, 5):
    run(v)
torch.cuda.synchronize()
for v in (3, 5):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        run(v)
    e1.record()
    e1.synchronize()
    us = e0.elapsed_time(e1) * 100
    print(f"conv_tc{v}: {us:.1f} us  {2.0 * N * h * h * co * ci * r * r / us / 1e6:.0f} TFLOP/s")
torch.cuda.profiler.start()
run(3)
run(5)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
