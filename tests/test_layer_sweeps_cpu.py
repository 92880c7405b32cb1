"""Wide parametrised sweeps of every layer type against torch autograd, on both CPU backends (ATen and this library's
native kernels): forward, input gradient, parameter gradients (accumulating), output-shape inference, config round trip
through the factory and clone.  The reference's per-layer test files (conv2d 36 cases, dense 26, maxpool 27, avgpool 21,
batchnorm 28, groupnorm 8) cover the same ground case by case; here the case lists are tables."""
import pytest
import torch
import torch.nn.functional as F

from dcnn_b200 import nn as dnn
from dcnn_b200.nn.factory import LayerFactory
from dcnn_b200.nn.layers import (ActivationLayer, AvgPool2DLayer, BatchNormLayer, Conv2DLayer, DenseLayer, DropoutLayer,
                                 FlattenLayer, GroupNormLayer, MaxPool2DLayer)

BACKENDS = ["aten", "native"]


@pytest.fixture(params=BACKENDS)
def backend(request, monkeypatch):
    monkeypatch.setenv("DCNN_CPU_BACKEND", request.param)
    torch.manual_seed(11)
    return request.param


def _rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-12))


def _contract(layer, in_shape, out):
    """Shape inference, config round trip and clone agree with the live layer."""
    assert list(layer.compute_output_shape(list(in_shape))) == list(out.shape)
    cfg = layer.get_config()
    again = LayerFactory.create(layer.type(), cfg)
    assert again.get_config() == cfg and type(again) is type(layer)
    assert layer.clone().get_config() == cfg


# ------------------------------------------------------------------------------------------------------ convolution
CONV = [  # N, Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, bias
    (2, 3, 8, 8, 4, 3, 3, 1, 1, 1, 1, True), (1, 1, 5, 5, 1, 1, 1, 1, 1, 0, 0, False), (2, 4, 9, 7, 6, 3, 3, 2, 2, 1, 1, True),
    (3, 2, 10, 10, 5, 5, 5, 1, 1, 2, 2, False), (2, 3, 16, 16, 8, 7, 7, 2, 2, 3, 3, True), (1, 6, 6, 6, 3, 1, 1, 2, 2, 0, 0, True),
    (2, 2, 7, 9, 4, 3, 5, 1, 2, 1, 2, True), (2, 5, 4, 4, 7, 4, 4, 1, 1, 0, 0, False), (4, 8, 6, 6, 16, 3, 3, 1, 1, 0, 0, True),
    (2, 3, 5, 11, 2, 5, 3, 2, 1, 2, 1, False), (1, 16, 4, 4, 32, 3, 3, 1, 1, 1, 1, True), (5, 1, 12, 12, 3, 2, 2, 2, 2, 0, 0, True),
]


@pytest.mark.parametrize("cfg", CONV, ids=[f"conv{i}" for i in range(len(CONV))])
def test_conv2d_sweep(backend, cfg):
    N, Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, bias = cfg
    layer = Conv2DLayer(Cin, Cout, kh, kw, sh, sw, ph, pw, bias, "c")
    layer.set_seed(1)
    layer.initialize()
    x = torch.randn(N, Cin, H, W)
    y = layer.forward(x)
    xt = x.clone().requires_grad_(True)
    wt = layer.weights.clone().requires_grad_(True)
    bt = layer.bias.reshape(-1).clone().requires_grad_(True) if bias else None
    yt = F.conv2d(xt, wt, bt, (sh, sw), (ph, pw))
    assert _rel(y, yt.detach()) < 1e-5
    g = torch.randn_like(y)
    layer.weight_grad.fill_(0.25)                        # gradients accumulate on top of what is there
    dx = layer.backward(g)
    yt.backward(g)
    assert _rel(dx, xt.grad) < 1e-5 and _rel(layer.weight_grad - 0.25, wt.grad) < 1e-4
    if bias:
        assert _rel(layer.bias_grad.reshape(-1), bt.grad) < 1e-5
    _contract(layer, x.shape, y)
    assert len(layer.parameters()) == (2 if bias else 1) == len(layer.gradients())


# ------------------------------------------------------------------------------------------------------------ dense
DENSE = [(1, 1, 1, True), (4, 7, 3, True), (2, 32, 10, False), (8, 100, 1, True), (3, 5, 64, False), (16, 192, 10, True)]


@pytest.mark.parametrize("cfg", DENSE, ids=[f"dense{i}" for i in range(len(DENSE))])
def test_dense_sweep(backend, cfg):
    N, fin, fout, bias = cfg
    layer = DenseLayer(fin, fout, bias, "d")
    layer.set_seed(2)
    layer.initialize()
    x = torch.randn(N, fin, 1, 1)
    y = layer.forward(x)
    xt = x.reshape(N, fin).clone().requires_grad_(True)
    wt = layer.weights.reshape(fout, fin).clone().requires_grad_(True)
    bt = layer.bias.reshape(-1).clone().requires_grad_(True) if bias else None
    yt = F.linear(xt, wt, bt)
    assert y.shape == (N, fout, 1, 1) and _rel(y.reshape(N, fout), yt.detach()) < 1e-5
    g = torch.randn_like(y)
    dx = layer.backward(g)
    yt.backward(g.reshape(N, fout))
    assert _rel(dx.reshape(N, fin), xt.grad) < 1e-5 and _rel(layer.weight_grad.reshape(fout, fin), wt.grad) < 1e-5
    if bias:
        assert _rel(layer.bias_grad.reshape(-1), bt.grad) < 1e-5
    _contract(layer, x.shape, y)


# ---------------------------------------------------------------------------------------------------------- pooling
POOL = [  # N, C, H, W, kh, kw, sh, sw, ph, pw
    (2, 3, 8, 8, 2, 2, 2, 2, 0, 0), (1, 1, 5, 5, 3, 3, 1, 1, 1, 1), (2, 4, 9, 7, 3, 3, 2, 2, 1, 1), (2, 2, 6, 6, 2, 3, 2, 3, 0, 1),
    (3, 5, 4, 4, 4, 4, 4, 4, 0, 0), (1, 8, 24, 24, 3, 3, 3, 3, 0, 0), (2, 3, 7, 7, 3, 3, 2, 2, 0, 0), (2, 2, 5, 9, 2, 2, 1, 1, 1, 1),
]


@pytest.mark.parametrize("cfg", POOL, ids=[f"pool{i}" for i in range(len(POOL))])
@pytest.mark.parametrize("kind", ["max", "avg"])
def test_pool_sweep(backend, kind, cfg):
    N, C, H, W, kh, kw, sh, sw, ph, pw = cfg
    layer = (MaxPool2DLayer if kind == "max" else AvgPool2DLayer)(kh, kw, sh, sw, ph, pw, "p")
    x = torch.randn(N, C, H, W)
    y = layer.forward(x)
    xt = x.clone().requires_grad_(True)
    yt = F.max_pool2d(xt, (kh, kw), (sh, sw), (ph, pw)) if kind == "max" \
        else F.avg_pool2d(xt, (kh, kw), (sh, sw), (ph, pw), count_include_pad=True)
    assert _rel(y, yt.detach()) < 1e-6
    g = torch.randn_like(y)
    dx = layer.backward(g)
    yt.backward(g)
    assert _rel(dx, xt.grad) < 1e-6
    _contract(layer, x.shape, y)
    assert layer.parameters() == [] and not layer.has_parameters()


# ---------------------------------------------------------------------------------------------------- normalisation
BN = [(4, 3, 5, 5, True), (2, 8, 1, 1, True), (8, 1, 7, 3, False), (3, 16, 4, 4, True), (16, 4, 2, 2, False), (2, 32, 6, 6, True)]


@pytest.mark.parametrize("cfg", BN, ids=[f"bn{i}" for i in range(len(BN))])
def test_batchnorm_sweep(backend, cfg):
    N, C, H, W, affine = cfg
    layer = BatchNormLayer(C, 1e-5, 0.1, affine, "bn")
    layer.initialize()
    ref = torch.nn.BatchNorm2d(C, 1e-5, 0.1, affine)
    if affine:
        with torch.no_grad():
            layer.gamma.copy_(torch.rand(C, 1, 1, 1) + 0.5)
            layer.beta.copy_(torch.randn(C, 1, 1, 1))
            ref.weight.copy_(layer.gamma.reshape(-1))
            ref.bias.copy_(layer.beta.reshape(-1))
    x = torch.randn(N, C, H, W) * 3 + 1
    y = layer.forward(x)
    xt = x.clone().requires_grad_(True)
    yt = ref(xt)
    assert _rel(y, yt.detach()) < 1e-4
    g = torch.randn_like(y)
    dx = layer.backward(g)
    yt.backward(g)
    # tiny per-channel sample counts amplify rounding, and with two samples per channel the gradient is pure rounding
    assert float((dx - xt.grad).norm()) < 2e-3 * float(xt.grad.norm()) + 1e-4 * float(g.norm())
    if affine:
        assert _rel(layer.gamma_grad.reshape(-1), ref.weight.grad) < 1e-4
        assert _rel(layer.beta_grad.reshape(-1), ref.bias.grad) < 1e-4
    assert _rel(layer.running_mean, ref.running_mean) < 1e-4 and _rel(layer.running_var, ref.running_var) < 1e-4
    layer.set_training(False)
    ref.eval()
    assert _rel(layer.forward(x), ref(x).detach()) < 1e-4
    _contract(layer, x.shape, y)


GN = [(2, 4, 2, 5, 5, True), (3, 8, 8, 3, 3, True), (1, 6, 1, 4, 4, False), (4, 12, 3, 2, 2, True)]


@pytest.mark.parametrize("cfg", GN, ids=[f"gn{i}" for i in range(len(GN))])
def test_groupnorm_sweep(backend, cfg):
    N, C, G, H, W, affine = cfg
    layer = GroupNormLayer(G, C, 1e-5, affine, "gn")
    layer.initialize()
    ref = torch.nn.GroupNorm(G, C, 1e-5, affine)
    if affine:
        with torch.no_grad():
            layer.gamma.copy_(torch.rand(C, 1, 1, 1) + 0.5)
            layer.beta.copy_(torch.randn(C, 1, 1, 1))
            ref.weight.copy_(layer.gamma.reshape(-1))
            ref.bias.copy_(layer.beta.reshape(-1))
    x = torch.randn(N, C, H, W) * 2 - 1
    y = layer.forward(x)
    xt = x.clone().requires_grad_(True)
    yt = ref(xt)
    assert _rel(y, yt.detach()) < 1e-4
    g = torch.randn_like(y)
    dx = layer.backward(g)
    yt.backward(g)
    assert _rel(dx, xt.grad) < 1e-3
    if affine:
        assert _rel(layer.gamma_grad.reshape(-1), ref.weight.grad) < 1e-4
        assert _rel(layer.beta_grad.reshape(-1), ref.bias.grad) < 1e-4
    _contract(layer, x.shape, y)


# ------------------------------------------------------------------------------------------ activations and the rest
ACT = [("relu", F.relu), ("leaky_relu", lambda t: F.leaky_relu(t, 0.01)), ("elu", F.elu), ("sigmoid", torch.sigmoid),
       ("tanh", torch.tanh), ("softmax", lambda t: torch.softmax(t, dim=1)), ("linear", lambda t: t)]


@pytest.mark.parametrize("name,fn", ACT, ids=[a[0] for a in ACT])
@pytest.mark.parametrize("shape", [(2, 3, 4, 4), (5, 10, 1, 1), (1, 1, 1, 7)])
def test_activation_sweep(backend, name, fn, shape):
    layer = ActivationLayer(name, "a")
    x = torch.randn(*shape)
    y = layer.forward(x)
    xt = x.clone().requires_grad_(True)
    yt = fn(xt)
    assert _rel(y, yt.detach()) < 1e-5
    g = torch.randn_like(y)
    dx = layer.backward(g)
    yt.backward(g)
    assert _rel(dx, xt.grad) < 1e-4
    _contract(layer, x.shape, y)


@pytest.mark.parametrize("rate", [0.0, 0.25, 0.5, 0.9])
def test_dropout_sweep(backend, rate):
    layer = DropoutLayer(rate, "do")
    layer.set_seed(4)
    x = torch.ones(64, 8, 4, 4)
    y = layer.forward(x)
    kept = (y != 0).float().mean().item()
    assert abs(kept - (1 - rate)) < 0.05
    if rate > 0:
        assert torch.allclose(y[y != 0], torch.full_like(y[y != 0], 1.0 / (1 - rate)))
    g = torch.ones_like(y)
    dx = layer.backward(g)
    assert torch.equal(dx != 0, y != 0)                  # the same mask is regenerated in backward
    layer.set_training(False)
    assert torch.equal(layer.forward(x), x)
    _contract(layer, x.shape, y)


@pytest.mark.parametrize("shape", [(2, 3, 4, 5), (1, 1, 1, 1), (7, 48, 2, 2)])
def test_flatten_sweep(backend, shape):
    layer = FlattenLayer("f")
    x = torch.randn(*shape)
    y = layer.forward(x)
    assert y.shape == (shape[0], shape[1] * shape[2] * shape[3], 1, 1) and torch.equal(y.reshape(shape), x)
    assert torch.equal(layer.backward(y), x)
    _contract(layer, x.shape, y)


@pytest.mark.parametrize("steps", [1, 3])
def test_repeated_forward_is_deterministic_and_buffers_do_not_leak(backend, steps):
    """The reference's layer_buffer_reuse_test: the same input gives the same output across calls and epochs."""
    m = (dnn.SequentialBuilder("reuse").input([3, 10, 10]).conv2d(4, 3, 3, 1, 1, 1, 1).batchnorm().activation("relu")
         .maxpool2d(2, 2, 2, 2).basic_residual_block(4, 8, 2).avgpool2d(3, 3).flatten().dense(5).build())
    m.set_seed(9)
    m.initialize()
    m.set_training(False)
    x = torch.randn(4, 3, 10, 10)
    first = m.forward(x).clone()
    for _ in range(steps):
        m.forward(torch.randn(2, 3, 10, 10))             # another shape in between must not disturb anything
        assert torch.equal(m.forward(x), first)
