"""ncu target: two EAGER training steps of the headline model (ResNet-50 64x64, batch 256, Adam) -- the first warms
everything up, the second runs between cudaProfilerStart/Stop so that `ncu --profile-from-start off -k regex:<kernel>
-c 1` captures the first launch of that kernel in a steady-state step.  Run it plain first (it must exit 0)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200 import nn as dnn  # noqa: E402
from dcnn_b200.models import create_resnet50_tiny_imagenet  # noqa: E402
from dcnn_b200.train import TrainStep  # noqa: E402

B = int(os.environ.get("NCU_BATCH", "256"))
m = create_resnet50_tiny_imagenet()
m.set_seed(0)
m.set_device("cuda:0")
m.initialize()
m.compute_input_gradient = False
opt = dnn.Adam(1e-4, epsilon=1e-3, weight_decay=1e-3)
opt.attach(m.parameters(), m.gradients())
step = TrainStep(m, opt, dnn.LogSoftmaxCrossEntropyLoss(), use_cuda_graph=False)
x = torch.randn(B, 3, 64, 64, device="cuda:0")
t = torch.nn.functional.one_hot(torch.randint(0, 200, (B,), device="cuda:0"), 200).float().reshape(B, 200, 1, 1)
step(x, t)
torch.cuda.synchronize()
torch.cuda.profiler.start()
loss, _ = step(x, t)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("loss", float(loss))
