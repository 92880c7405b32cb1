"""ncu target: the weight-gradient kernel on full-size ResNet-50 layers (batch 256): 3x3 256->256 @16x16 (wgrad<128,3>)
and 3x3 64->64 @32x32 (wgrad<64,4>, two taps per CTA), one launch each between cudaProfilerStart/Stop."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200.ops import cuda as K  # noqa: E402

CL = torch.channels_last
N = 256
cases = []
for ci, co, h in ((256, 256, 16), (64, 64, 32)):
    x = torch.randn(N, ci, h, h, device="cuda").contiguous(memory_format=CL)
    dy = torch.randn(N, co, h, h, device="cuda").contiguous(memory_format=CL)
    dw = torch.zeros(co, ci, 3, 3, device="cuda").contiguous(memory_format=CL)
    cases.append((f"3x3 {ci}->{co} @{h}", x, dy, dw))

for name, x, dy, dw in cases:
    for _ in range(3):
        K.conv2d_backward_weight(dy, x, dw, (1, 1), (1, 1))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        K.conv2d_backward_weight(dy, x, dw, (1, 1), (1, 1))
    e1.record()
    e1.synchronize()
    print(f"wgrad {name}: {e0.elapsed_time(e1) * 100:.1f} us")
torch.cuda.profiler.start()
for name, x, dy, dw in cases:
    K.conv2d_backward_weight(dy, x, dw, (1, 1), (1, 1))
torch.cuda.synchronize()
torch.cuda.profiler.stop()
