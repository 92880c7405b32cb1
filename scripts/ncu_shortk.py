"""ncu target: an epilogue-bound convolution (1x1 64->256 @32x32, batch 256, with BatchNorm statistics) and the
BatchNorm backward reduce on a 268 MB tensor (256 channels @32x32), one launch each between cudaProfilerStart/Stop."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200 import ops  # noqa: E402
from dcnn_b200.ops import cuda as K  # noqa: E402

CL = torch.channels_last
N = 256
x = torch.randn(N, 64, 32, 32, device="cuda").contiguous(memory_format=CL)
w = (torch.randn(256, 64, 1, 1, device="cuda") * 0.1).contiguous(memory_format=CL)
gamma, beta = torch.ones(256, 1, 1, 1, device="cuda"), torch.zeros(256, 1, 1, 1, device="cuda")


def step():
    y, st = K.conv2d_forward(x, w, None, (1, 1), (0, 0), want_stats=True)
    rm, rv = torch.zeros(256, device="cuda"), torch.ones(256, device="cuda")
    mean, inv = ops.bn_finalize(st, 1e-5, 0.1, rm, rv, gamma=gamma, beta=beta)
    res = torch.randn_like(y)
    mask = torch.zeros(y.numel() // 4, dtype=torch.uint8, device="cuda")
    out = ops.bn_apply(y, mean, inv, gamma, beta, res, True, mask_out=mask)
    dg, db = torch.zeros_like(gamma), torch.zeros_like(beta)
    dx, m = ops.bn_backward(torch.randn_like(y), y, None, mean, inv, gamma, dg, db, True, True, relu_mask=mask)
    return dx


for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(10):
    K.conv2d_forward(x, w, None, (1, 1), (0, 0), want_stats=True)
e1.record()
e1.synchronize()
print(f"conv 1x1 64->256 @32 + stats: {e0.elapsed_time(e1) * 100:.1f} us")
torch.cuda.profiler.start()
step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
