"""Small driver for profiling: a few eager training steps of ResNet-50 64x64 (batch from argv)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200 import nn
from dcnn_b200.models import create_resnet50_tiny_imagenet
from dcnn_b200.train import TrainStep
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
torch.manual_seed(0)
m = create_resnet50_tiny_imagenet(); m.set_seed(1); m.set_device("cuda:0"); m.initialize(); m.compute_input_gradient = False
opt = nn.Adam(1e-4, epsilon=1e-3, weight_decay=1e-3); opt.attach(m.parameters(), m.gradients())
step = TrainStep(m, opt, nn.LogSoftmaxCrossEntropyLoss())
x = torch.randn(B, 3, 64, 64, device="cuda"); t = torch.nn.functional.one_hot(torch.randint(0, 200, (B,)), 200).float().reshape(B, 200, 1, 1).cuda()
for i in range(steps):
    loss, c = step(x, t)
torch.cuda.synchronize()
print("loss", float(loss))
