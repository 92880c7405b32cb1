"""Sequential -> torch.nn.Module export: the exported module computes the same function (outputs, input and parameter
gradients, BatchNorm statistics updates) for every layer family and for zoo models, and the weights round-trip."""
import pytest
import torch

from dcnn_b200 import nn
from dcnn_b200.io import load_from_torch_module, to_torch_module
from dcnn_b200.models import MODEL_ZOO


def _every_layer_model():
    b = nn.SequentialBuilder("all_layers").input([3, 16, 16])
    b.conv2d(8, 3, 3, 1, 1, 1, 1, True, "c1").batchnorm(1e-5, 0.1, True, "bn1").activation("leaky_relu", "a1")
    b.basic_residual_block(8, 8, 1, "rb_identity").basic_residual_block(8, 16, 2, "rb_projection")
    b.conv2d(16, 1, 1, 1, 1, 0, 0, False, "c2").groupnorm(4, 1e-5, True, "gn").activation("elu", "a2")
    b.maxpool2d(3, 3, 2, 2, 1, 1, "mp").avgpool2d(2, 2, 2, 2, 0, 0, "ap").activation("tanh", "a3")
    b.flatten("fl").dense(12, True, "d1").activation("sigmoid", "a4").dense(5, False, "d2")
    return b.build()


def _rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-12))


# the tolerance for the 50-layer network covers fp32 rounding amplified through 53 BatchNorms over 4-sample batches
@pytest.mark.parametrize("make,shape,tol", [(_every_layer_model, (4, 3, 16, 16), 1e-4),
                                            (MODEL_ZOO["mnist_cnn"], (4, 1, 28, 28), 1e-4),
                                            (MODEL_ZOO["resnet9_cifar10"], (2, 3, 32, 32), 1e-3),
                                            (MODEL_ZOO["resnet50_tiny_imagenet"], (4, 3, 64, 64), 3e-2)])
def test_exported_module_matches(make, shape, tol):
    torch.manual_seed(0)
    model = make()
    model.set_seed(5)
    model.initialize()
    model.set_training(True)
    net = to_torch_module(model)
    assert net.training
    x = torch.randn(*shape)
    y = model.forward(x)
    xt = x.clone().requires_grad_(True)
    yt = net(xt)
    assert y.shape == yt.shape and _rel(y, yt.detach()) < tol
    g = torch.randn_like(y)
    dx = model.backward(g)
    yt.backward(g)
    assert _rel(dx, xt.grad) < tol
    ours = [t for t in model.gradients()]
    theirs = [p.grad for p in net.parameters()]
    assert len(ours) == len(theirs)
    scale = max(float(b.norm()) for b in theirs)
    for a, b in zip(ours, theirs):
        # (a bias feeding a BatchNorm has an exactly-zero gradient: both sides hold rounding noise there, of a size
        # that follows the other gradients' magnitude)
        assert float((a.reshape(b.shape) - b).norm()) < 10 * tol * float(b.norm()) + 1e-4 * scale
    # BatchNorm running statistics moved the same way; evaluation mode agrees
    model.set_training(False)
    net.eval()
    with torch.no_grad():
        assert _rel(model.forward(x), net(x)) < tol


def test_weights_round_trip_from_torch():
    model = MODEL_ZOO["resnet9_cifar10"]()
    model.set_seed(1)
    model.initialize()
    net = to_torch_module(model)
    with torch.no_grad():
        for p in net.parameters():
            p.add_(0.01)
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.add_(0.5)
    load_from_torch_module(model, net)
    model.set_training(False)
    net.eval()
    x = torch.randn(2, 3, 32, 32)
    with torch.no_grad():
        assert torch.allclose(model.forward(x), net(x), atol=2e-4, rtol=1e-3)
