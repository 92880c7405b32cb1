"""Per-kernel device-time shares of one eager ResNet-50 (64x64, batch 256, Adam) training step, via CUPTI activity
records (torch.profiler, CUDA activities only).  Shares only -- absolute step time comes from bench.py, never from a
profiled run."""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200 import nn  # noqa: E402
from dcnn_b200.models import create_resnet50_tiny_imagenet  # noqa: E402
from dcnn_b200.train import TrainStep  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = 2
torch.manual_seed(0)
m = create_resnet50_tiny_imagenet()
m.set_seed(1)
m.set_device("cuda:0")
m.initialize()
m.compute_input_gradient = False
opt = nn.Adam(1e-4, epsilon=1e-3, weight_decay=1e-3)
opt.attach(m.parameters(), m.gradients())
step = TrainStep(m, opt, nn.LogSoftmaxCrossEntropyLoss())
x = torch.randn(B, 3, 64, 64, device="cuda")
t = torch.nn.functional.one_hot(torch.randint(0, 200, (B,)), 200).float().reshape(B, 200, 1, 1).cuda()
for _ in range(2):
    step(x, t)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(steps):
        step(x, t)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0.0, 0])
for ev in prof.events():
    dt = getattr(ev, "device_time_total", None)
    if dt is None:
        dt = getattr(ev, "cuda_time_total", 0.0)
    if dt and dt > 0:
        agg[ev.name][0] += dt
        agg[ev.name][1] += 1
tot = sum(v[0] for v in agg.values())
print(f"kernel time per step: {tot / steps / 1e3:.3f} ms over {sum(v[1] for v in agg.values()) // steps} launches "
      f"(DCNN_CONV_V3={os.environ.get('DCNN_CONV_V3', '0')})")
print(f"{'share':>6s} {'ms/step':>8s} {'n/step':>6s} {'avg us':>8s}  kernel")
for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
    print(f"{100 * us / tot:5.1f}% {us / steps / 1e3:8.3f} {n // steps:6d} {us / n:8.1f}  {name[:110]}")
