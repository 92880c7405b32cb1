"""CPU layer tests.  Oracle = torch autograd on the same maths (the reference's layer tests use naive loops
written inside the test, unit_tests/conv2d_layer_test.cpp:65-123; autograd is the equivalent independent check)."""
import json

import pytest
import torch
import torch.nn.functional as F

import dcnn_b200 as dc
from dcnn_b200 import nn as dnn
from dcnn_b200.models import MODEL_ZOO, create_mnist_trainer, create_resnet9_cifar10

torch.manual_seed(0)


def _autograd_check(layer, x, atol=1e-4, torch_fn=None):
    layer.set_seed(3)
    layer.initialize()
    y = layer.forward(x, 0)
    g = torch.randn_like(y)
    dx = layer.backward(g, 0)
    xr = x.clone().requires_grad_(True)
    params = [p.clone().requires_grad_(True) for p in layer.parameters()]
    yr = torch_fn(xr, params)
    assert torch.allclose(y, yr, atol=atol, rtol=1e-4), (y - yr).abs().max()
    grads = torch.autograd.grad(yr, [xr] + params, g)
    assert torch.allclose(dx, grads[0], atol=atol, rtol=1e-4)
    for mine, ref in zip(layer.gradients(), grads[1:]):
        assert torch.allclose(mine, ref.reshape(mine.shape), atol=atol, rtol=1e-3)


@pytest.mark.parametrize("k,s,p,bias", [(3, 1, 1, True), (1, 1, 0, False), (3, 2, 1, True), (5, 1, 0, True), (7, 2, 3, False)])
def test_conv2d(k, s, p, bias):
    l = dnn.Conv2DLayer(3, 8, k, k, s, s, p, p, bias, "c")
    _autograd_check(l, torch.randn(2, 3, 13, 13),
                    torch_fn=lambda x, P: F.conv2d(x, P[0], P[1].reshape(-1) if bias else None, stride=s, padding=p))


def test_dense():
    l = dnn.DenseLayer(12, 7, True, "d")
    _autograd_check(l, torch.randn(5, 3, 2, 2),
                    torch_fn=lambda x, P: (x.reshape(5, -1) @ P[0].reshape(7, 12).t() + P[1].reshape(1, -1)).reshape(5, 7, 1, 1))


def test_batchnorm_train_and_running_stats():
    l = dnn.BatchNormLayer(6, 1e-5, 0.1, True, "bn")
    x = torch.randn(4, 6, 5, 5) * 2 + 1
    _autograd_check(l, x, torch_fn=lambda x, P: F.batch_norm(x, None, None, P[0].reshape(-1), P[1].reshape(-1), True, 0.1, 1e-5))
    ref = torch.nn.BatchNorm2d(6, eps=1e-5, momentum=0.1)
    ref.train()
    ref(x)
    assert torch.allclose(l.running_mean, ref.running_mean, atol=1e-5)
    assert torch.allclose(l.running_var, ref.running_var, atol=1e-5)
    l.set_training(False)
    ref.eval()
    assert torch.allclose(l.forward(x), ref(x), atol=1e-4)


def test_groupnorm():
    l = dnn.GroupNormLayer(2, 6, 1e-5, True, "gn")
    _autograd_check(l, torch.randn(3, 6, 4, 4),
                    torch_fn=lambda x, P: F.group_norm(x, 2, P[0].reshape(-1), P[1].reshape(-1), 1e-5))


@pytest.mark.parametrize("k,s,p", [(2, 2, 0), (3, 2, 1), (3, 3, 0), (3, 1, 1)])
def test_maxpool(k, s, p):
    l = dnn.MaxPool2DLayer(k, k, s, s, p, p, "mp")
    _autograd_check(l, torch.randn(2, 3, 9, 9), torch_fn=lambda x, P: F.max_pool2d(x, k, s, p))


@pytest.mark.parametrize("k,s,p", [(2, 2, 0), (4, 1, 0), (3, 2, 1)])
def test_avgpool(k, s, p):
    l = dnn.AvgPool2DLayer(k, k, s, s, p, p, "ap")
    _autograd_check(l, torch.randn(2, 3, 8, 8), torch_fn=lambda x, P: F.avg_pool2d(x, k, s, p, count_include_pad=True))


@pytest.mark.parametrize("name,fn", [("relu", torch.relu), ("leaky_relu", lambda x: F.leaky_relu(x, 0.01)),
                                     ("elu", F.elu), ("sigmoid", torch.sigmoid), ("tanh", torch.tanh),
                                     ("softmax", lambda x: torch.softmax(x, 1)), ("linear", lambda x: x * 1)])
def test_activations(name, fn):
    l = dnn.ActivationLayer(name, "act")
    _autograd_check(l, torch.randn(3, 5, 4, 4), torch_fn=lambda x, P: fn(x))


def test_dropout_mask_consistency():
    l = dnn.DropoutLayer(0.5, "do")
    l.set_seed(7)
    x = torch.ones(4, 3, 8, 8)
    y = l.forward(x, 0)
    assert set(y.unique().tolist()) <= {0.0, 2.0}
    g = l.backward(torch.ones_like(x), 0)
    assert torch.equal(g, y)
    l.set_training(False)
    assert torch.equal(l.forward(x), x)


def test_flatten_roundtrip():
    l = dnn.FlattenLayer("f")
    x = torch.randn(2, 3, 4, 5)
    y = l.forward(x)
    assert y.shape == (2, 60, 1, 1)
    assert torch.equal(l.backward(y), x)


def _model_autograd(model, x, fuse):
    """Run our explicit fwd/bwd and compare every parameter gradient with autograd through torch ops."""
    model.fuse = fuse
    model._units = None
    model.clear_gradients()
    y = model.forward(x)
    g = torch.randn_like(y)
    dx = model.backward(g)
    return y, dx, [t.clone() for t in model.gradients()], g


def test_fused_equals_unfused_resnet9():
    m = create_resnet9_cifar10()
    m.set_seed(11)
    m.initialize()
    x = torch.randn(4, 3, 32, 32)
    y1, dx1, g1, g = _model_autograd(m, x, True)
    m.fuse = False
    m._units = None
    m.clear_gradients()
    # reset running stats effect is irrelevant for train-mode outputs
    y2 = m.forward(x)
    dx2 = m.backward(g)
    assert torch.allclose(y1, y2, atol=1e-5)
    assert torch.allclose(dx1, dx2, atol=1e-5)
    for a, b in zip(g1, m.gradients()):
        assert torch.allclose(a, b, atol=1e-4, rtol=1e-3)


def test_residual_block_vs_autograd():
    b = dnn.SequentialBuilder("t").input([8, 6, 6]).bottleneck_residual_block(8, 4, 16, 2, "blk").build()
    b.set_seed(5)
    b.initialize()
    blk = b.layers[0]
    x = torch.randn(3, 8, 6, 6)
    y = b.forward(x)
    g = torch.randn_like(y)
    dx = b.backward(g)

    P = [p.clone().requires_grad_(True) for p in blk.parameters()]
    xr = x.clone().requires_grad_(True)

    def bn(t, gamma, beta):
        return F.batch_norm(t, None, None, gamma.reshape(-1), beta.reshape(-1), True, 0.1, 1e-3)
    h = torch.relu(bn(F.conv2d(xr, P[0]), P[1], P[2]))
    h = torch.relu(bn(F.conv2d(h, P[3], stride=2, padding=1), P[4], P[5]))
    h = bn(F.conv2d(h, P[6]), P[7], P[8])
    sc = bn(F.conv2d(xr, P[9], stride=2), P[10], P[11])
    yr = torch.relu(h + sc)
    assert torch.allclose(y, yr, atol=1e-4)
    grads = torch.autograd.grad(yr, [xr] + P, g)
    assert torch.allclose(dx, grads[0], atol=1e-4, rtol=1e-3)
    for mine, ref in zip(blk.gradients(), grads[1:]):
        assert torch.allclose(mine, ref.reshape(mine.shape), atol=1e-4, rtol=1e-3)


def test_microbatch_gradients_equal_sum_of_microbatches():
    """Weight gradients accumulate across microbatches (beta=1 semantics) and caches are per microbatch."""
    m = create_mnist_trainer()
    m.set_seed(2)
    m.initialize()
    x = torch.randn(8, 1, 28, 28)
    ys = [m.forward(x[i * 4:(i + 1) * 4], i) for i in range(2)]
    gs = [torch.randn_like(y) for y in ys]
    for i in (1, 0):
        m.backward(gs[i], i)
    acc = [g.clone() for g in m.gradients()]
    total = [torch.zeros_like(g) for g in acc]
    for i in range(2):
        m.clear_gradients()
        m.forward(x[i * 4:(i + 1) * 4], 0)
        m.backward(gs[i], 0)
        for t, g in zip(total, m.gradients()):
            t.add_(g)
    for a, b in zip(acc, total):
        assert torch.allclose(a, b, atol=1e-5, rtol=1e-4)


@pytest.mark.parametrize("name", sorted(MODEL_ZOO))
def test_zoo_builds_and_shapes(name):
    m = MODEL_ZOO[name]()
    cfg = m.get_config()
    first = cfg["layers"][0]["parameters"]
    hw = {"mnist_cnn": 28, "resnet50_imagenet": 224}.get(name, 64 if "tiny" in name else 32)
    out = m.compute_output_shape([2, first["in_channels"], hw, hw])
    classes = {"mnist_cnn": 10, "resnet50_imagenet": 1000, "cifar100_cnn": 100}.get(name, 200 if "tiny" in name else 10)
    assert out == [2, classes, 1, 1]
    m2 = dc.Sequential.load_from_config(json.loads(json.dumps(cfg)))
    assert m2.get_config() == cfg
