"""End-to-end GPU tests: whole models through the public API on cuda:0 against the CPU path."""
import pytest
import torch

import dcnn_b200 as dc
from dcnn_b200 import nn as dnn
from dcnn_b200.models import create_mnist_trainer, create_resnet9_cifar10, create_resnet50_tiny_imagenet

pytestmark = pytest.mark.gpu


def _clone_to(model, device):
    m = dc.Sequential.load_from_config(model.get_config())
    m.initialize()
    m.load_parameters([p.detach().cpu() for p in model.parameters()])
    m.set_device(device)
    return m


def _rel(a, b):
    a, b = a.detach().cpu().float(), b.detach().cpu().float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


@pytest.mark.parametrize("factory,shape,tol", [
    (create_mnist_trainer, (8, 1, 28, 28), 5e-4),
    (create_resnet9_cifar10, (8, 3, 32, 32), 3e-2),
])
def test_model_forward_backward_matches_cpu(factory, shape, tol):
    torch.manual_seed(0)
    cpu = factory()
    cpu.set_seed(1)
    cpu.initialize()
    gpu = _clone_to(cpu, "cuda:0")
    x = torch.randn(*shape)
    y_c = cpu.forward(x)
    y_g = gpu.forward(x.cuda())
    assert _rel(y_g, y_c) < tol
    g = torch.randn_like(y_c)
    dx_c = cpu.backward(g)
    dx_g = gpu.backward(g.cuda())
    assert _rel(dx_g, dx_c) < 10 * tol
    worst = max(_rel(a, b) for a, b in zip(gpu.gradients(), cpu.gradients()))
    assert worst < 20 * tol, worst


def test_resnet50_bottleneck_stage_matches_cpu():
    torch.manual_seed(0)
    b = (dnn.SequentialBuilder("t").input([64, 16, 16])
         .bottleneck_residual_block(64, 64, 256, 1, "b1")
         .bottleneck_residual_block(256, 128, 512, 2, "b2")
         .bottleneck_residual_block(512, 128, 512, 1, "b3").build())
    b.set_seed(3)
    b.initialize()
    g = _clone_to(b, "cuda:0")
    x = torch.randn(8, 64, 16, 16)
    yc, yg = b.forward(x), g.forward(x.cuda())
    assert _rel(yg, yc) < 2e-2
    gr = torch.randn_like(yc)
    dc_, dg_ = b.backward(gr), g.backward(gr.cuda())
    assert _rel(dg_, dc_) < 5e-2
    worst = max(_rel(a, bb) for a, bb in zip(g.gradients(), b.gradients()))
    assert worst < 8e-2, worst


def test_resnet50_training_step_runs_and_loss_decreases():
    torch.manual_seed(0)
    m = create_resnet50_tiny_imagenet()
    m.set_seed(1)
    m.set_device("cuda:0")
    m.initialize()
    m.compute_input_gradient = False
    opt = dnn.Adam(1e-3)
    opt.attach(m.parameters(), m.gradients())
    loss_fn = dnn.LogSoftmaxCrossEntropyLoss()
    x = torch.randn(16, 3, 64, 64, device="cuda")
    t = torch.nn.functional.one_hot(torch.randint(0, 200, (16,)), 200).float().reshape(16, 200, 1, 1).cuda()
    losses = []
    for _ in range(8):
        y = m.forward(x)
        loss, grad, correct = loss_fn.forward_backward(y, t)
        m.backward(grad)
        opt.update(zero_grads=True)
        losses.append(float(loss))
    assert all(l == l for l in losses), losses
    assert losses[-1] < losses[0], losses
