"""Layer-level contract tests on CPU, in the spirit of the reference's unit_tests/*_layer_test.cpp: a naive nested-loop
oracle written inside the test (independent of ATen), output-shape inference, config round-trip through the factory,
clone independence, repeated-forward determinism (the reference's layer_buffer_reuse_test.cpp), edge cases (1x1,
asymmetric kernels / strides / padding, 7x7 stride 2, batch 1), parameter / FLOP accounting and error paths."""
import json

import pytest
import torch

from dcnn_b200 import nn as dnn
from dcnn_b200.nn.factory import LayerFactory

torch.manual_seed(0)


# ------------------------------------------------------------------------------------------------ naive oracles
def naive_conv(x, w, b, sh, sw, ph, pw):
    N, C, H, W = x.shape
    OC, _, KH, KW = w.shape
    OH, OW = (H + 2 * ph - KH) // sh + 1, (W + 2 * pw - KW) // sw + 1
    y = torch.zeros(N, OC, OH, OW, dtype=torch.float64)
    for n in range(N):
        for oc in range(OC):
            for oh in range(OH):
                for ow in range(OW):
                    acc = float(b[oc]) if b is not None else 0.0
                    for c in range(C):
                        for r in range(KH):
                            for s in range(KW):
                                ih, iw = oh * sh - ph + r, ow * sw - pw + s
                                if 0 <= ih < H and 0 <= iw < W:
                                    acc += float(x[n, c, ih, iw]) * float(w[oc, c, r, s])
                    y[n, oc, oh, ow] = acc
    return y


def naive_conv_backward(x, w, g, sh, sw, ph, pw):
    N, C, H, W = x.shape
    OC, _, KH, KW = w.shape
    dx = torch.zeros_like(x, dtype=torch.float64)
    dw = torch.zeros_like(w, dtype=torch.float64)
    for n in range(N):
        for oc in range(OC):
            for oh in range(g.shape[2]):
                for ow in range(g.shape[3]):
                    gv = float(g[n, oc, oh, ow])
                    for c in range(C):
                        for r in range(KH):
                            for s in range(KW):
                                ih, iw = oh * sh - ph + r, ow * sw - pw + s
                                if 0 <= ih < H and 0 <= iw < W:
                                    dx[n, c, ih, iw] += gv * float(w[oc, c, r, s])
                                    dw[oc, c, r, s] += gv * float(x[n, c, ih, iw])
    return dx, dw, g.double().sum(dim=(0, 2, 3))


def naive_pool(x, kh, kw, sh, sw, ph, pw, mode):
    N, C, H, W = x.shape
    OH, OW = (H + 2 * ph - kh) // sh + 1, (W + 2 * pw - kw) // sw + 1
    y = torch.zeros(N, C, OH, OW)
    for n in range(N):
        for c in range(C):
            for oh in range(OH):
                for ow in range(OW):
                    vals = []
                    for r in range(kh):
                        for s in range(kw):
                            ih, iw = oh * sh - ph + r, ow * sw - pw + s
                            if 0 <= ih < H and 0 <= iw < W:
                                vals.append(float(x[n, c, ih, iw]))
                            elif mode == "avg":
                                vals.append(0.0)            # zero padding counts in the average (count_include_pad)
                    y[n, c, oh, ow] = max(vals) if mode == "max" else sum(vals) / (kh * kw)
    return y


# ------------------------------------------------------------------------------------------------ convolution
CONV_EDGE = [  # in, out, kh, kw, sh, sw, ph, pw, bias, N, H, W
    (2, 3, 1, 1, 1, 1, 0, 0, True, 2, 4, 5),
    (1, 2, 3, 3, 1, 1, 1, 1, False, 1, 5, 5),
    (2, 2, 3, 2, 2, 1, 1, 0, True, 2, 6, 5),        # asymmetric kernel / stride / padding
    (3, 4, 7, 7, 2, 2, 3, 3, True, 1, 9, 9),        # 7x7 stride 2 (ImageNet stem)
    (2, 2, 2, 3, 1, 2, 0, 1, False, 3, 4, 7),
]


@pytest.mark.parametrize("cfg", CONV_EDGE, ids=lambda c: "x".join(map(str, c)))
def test_conv2d_against_naive_loops(cfg):
    ci, co, kh, kw, sh, sw, ph, pw, bias, N, H, W = cfg
    layer = dnn.Conv2DLayer(ci, co, kh, kw, sh, sw, ph, pw, bias, "c")
    layer.set_seed(11)
    layer.initialize()
    x = torch.randn(N, ci, H, W)
    w = layer.parameters()[0]
    b = layer.parameters()[1].reshape(-1) if bias else None
    y = layer.forward(x, 0)
    want = naive_conv(x, w, b, sh, sw, ph, pw)
    assert list(y.shape) == layer.compute_output_shape([N, ci, H, W]) == list(want.shape)
    assert torch.allclose(y.double(), want, atol=1e-5)
    g = torch.randn_like(y)
    dx = layer.backward(g, 0)
    dxw, dww, dbw = naive_conv_backward(x, w, g, sh, sw, ph, pw)
    assert torch.allclose(dx.double(), dxw, atol=1e-5)
    assert torch.allclose(layer.gradients()[0].double(), dww, atol=1e-5)
    if bias:
        assert torch.allclose(layer.gradients()[1].double().reshape(-1), dbw, atol=1e-5)
    # gradients accumulate across backward calls until cleared (microbatching contract)
    layer.forward(x, 0)
    layer.backward(g, 0)
    assert torch.allclose(layer.gradients()[0].double(), 2 * dww, atol=1e-5)
    layer.clear_gradients()
    assert float(layer.gradients()[0].abs().sum()) == 0.0


def test_conv2d_accounting_and_errors():
    layer = dnn.Conv2DLayer(3, 8, 3, 3, 1, 1, 1, 1, True, "c")
    layer.initialize()
    assert sum(p.numel() for p in layer.parameters()) == 8 * 3 * 9 + 8
    assert layer.has_parameters() and layer.type() == "conv2d"
    assert layer.forward_flops([2, 3, 10, 10]) >= 2 * 2 * 8 * 3 * 9 * 100
    assert layer.backward_flops([2, 3, 10, 10]) >= layer.forward_flops([2, 3, 10, 10])
    with pytest.raises(Exception):
        layer.forward(torch.randn(2, 4, 10, 10), 0)      # wrong channel count
    with pytest.raises(Exception):
        layer.backward(torch.randn(2, 8, 10, 10), 5)     # no forward cached for this microbatch id


# ------------------------------------------------------------------------------------------------ pooling
@pytest.mark.parametrize("kh,kw,sh,sw,ph,pw", [(2, 2, 2, 2, 0, 0), (3, 3, 2, 2, 1, 1), (3, 2, 1, 2, 1, 0), (1, 1, 1, 1, 0, 0)])
def test_pools_against_naive_loops(kh, kw, sh, sw, ph, pw):
    x = torch.randn(2, 3, 7, 6)
    mp = dnn.MaxPool2DLayer(kh, kw, sh, sw, ph, pw, "mp")
    y = mp.forward(x, 0)
    assert torch.allclose(y, naive_pool(x, kh, kw, sh, sw, ph, pw, "max"), atol=1e-6)
    assert list(y.shape) == mp.compute_output_shape(list(x.shape))
    g = torch.randn_like(y)
    dx = mp.backward(g, 0)
    assert dx.shape == x.shape
    assert abs(float(dx.sum()) - float(g.sum())) < 1e-4      # each output routes its gradient to exactly one input
    ap = dnn.AvgPool2DLayer(kh, kw, sh, sw, ph, pw, "ap")
    ya = ap.forward(x, 0)
    assert torch.allclose(ya, naive_pool(x, kh, kw, sh, sw, ph, pw, "avg"), atol=1e-6)
    da = ap.backward(torch.ones_like(ya), 0)
    assert da.shape == x.shape and float(da.min()) >= 0.0


def test_maxpool_default_stride_equals_kernel():
    mp = dnn.MaxPool2DLayer(2, 2, name="mp")                 # stride 0 -> kernel size (reference default)
    assert mp.compute_output_shape([1, 4, 8, 8]) == [1, 4, 4, 4]
    assert mp.forward(torch.randn(1, 4, 8, 8), 0).shape == (1, 4, 4, 4)


# ------------------------------------------------------------------------------------------------ dense / norms
def test_dense_against_naive_and_4d_input():
    layer = dnn.DenseLayer(6, 4, True, "d")
    layer.set_seed(2)
    layer.initialize()
    x = torch.randn(3, 6, 1, 1)
    w, b = layer.parameters()[0].reshape(4, 6), layer.parameters()[1].reshape(4)
    y = layer.forward(x, 0)
    want = torch.tensor([[sum(float(x[n, k, 0, 0]) * float(w[o, k]) for k in range(6)) + float(b[o]) for o in range(4)]
                         for n in range(3)])
    assert torch.allclose(y.reshape(3, 4), want, atol=1e-5)
    assert layer.compute_output_shape([3, 6, 1, 1])[:2] == [3, 4]
    dx = layer.backward(torch.ones_like(y), 0)
    assert torch.allclose(dx.reshape(3, 6), w.sum(0).expand(3, 6), atol=1e-5)


def test_batchnorm_statistics_eval_mode_and_affine_off():
    bn = dnn.BatchNormLayer(3, 1e-5, 0.1, True, "bn")
    bn.initialize()
    x = torch.randn(16, 3, 4, 4) * 3 + 2
    y = bn.forward(x, 0)
    assert torch.allclose(y.mean(dim=(0, 2, 3)), torch.zeros(3), atol=1e-4)
    assert torch.allclose(y.var(dim=(0, 2, 3), unbiased=False), torch.ones(3), atol=1e-3)
    rm = bn.buffers()["running_mean"]
    assert torch.allclose(rm, 0.1 * x.mean(dim=(0, 2, 3)), atol=1e-4)
    n = x.numel() // 3
    assert torch.allclose(bn.buffers()["running_var"], 0.9 + 0.1 * x.var(dim=(0, 2, 3), unbiased=False) * n / (n - 1),
                          atol=1e-3)
    bn.set_training(False)
    ye = bn.forward(x, 0)
    want = (x - rm.view(1, 3, 1, 1)) / torch.sqrt(bn.buffers()["running_var"].view(1, 3, 1, 1) + 1e-5)
    assert torch.allclose(ye, want, atol=1e-4)
    plain = dnn.BatchNormLayer(3, 1e-5, 0.1, False, "bn2")
    plain.initialize()
    assert plain.parameters() == [] and not plain.has_parameters()
    assert plain.forward(x, 0).shape == x.shape
    with pytest.raises(ValueError):
        bn.forward(torch.randn(2, 5, 4, 4), 0)


def test_groupnorm_groups_normalise_independently():
    gn = dnn.GroupNormLayer(2, 4, 1e-5, True, "gn")
    gn.initialize()
    x = torch.randn(3, 4, 5, 5) * torch.tensor([1.0, 1.0, 10.0, 10.0]).view(1, 4, 1, 1)
    y = gn.forward(x, 0).reshape(3, 2, -1)
    assert torch.allclose(y.mean(-1), torch.zeros(3, 2), atol=1e-4)
    assert torch.allclose(y.var(-1, unbiased=False), torch.ones(3, 2), atol=1e-3)


# ------------------------------------------------------------------------------------------------ config / clone / reuse
ALL_LAYERS = [
    lambda: dnn.Conv2DLayer(3, 5, 3, 2, 2, 1, 1, 0, True, "conv"),
    lambda: dnn.DenseLayer(12, 7, False, "fc"),
    lambda: dnn.BatchNormLayer(3, 1e-4, 0.2, True, "bn"),
    lambda: dnn.GroupNormLayer(1, 3, 1e-5, True, "gn"),
    lambda: dnn.MaxPool2DLayer(3, 3, 2, 2, 1, 1, "mp"),
    lambda: dnn.AvgPool2DLayer(2, 2, 2, 2, 0, 0, "ap"),
    lambda: dnn.DropoutLayer(0.25, "do"),
    lambda: dnn.FlattenLayer("fl"),
    lambda: dnn.ActivationLayer(dnn.create_activation("leaky_relu"), "act"),
]


@pytest.mark.parametrize("make", ALL_LAYERS, ids=lambda m: m().type())
def test_config_roundtrip_clone_and_repeatability(make):
    layer = make()
    cfg = layer.get_config()
    json.loads(json.dumps(cfg))                                   # JSON-serialisable
    twin = LayerFactory.create(layer.type(), cfg)
    assert twin.type() == layer.type() and twin.get_config() == cfg and twin.name == layer.name
    clone = layer.clone()
    assert clone is not layer and clone.get_config() == cfg
    x = torch.randn(4, 12, 1, 1) if layer.type() == "dense" else torch.randn(4, 3, 8, 8)
    for l in (layer, twin):
        l.set_seed(5)
        l.initialize()
        l.set_training(False)
    y1, y2 = layer.forward(x, 0), twin.forward(x, 0)
    assert torch.equal(y1, y2)                                    # same seed + config -> same parameters
    assert list(y1.shape) == layer.compute_output_shape(list(x.shape))
    for _ in range(3):                                            # repeated forwards reuse caches without drift
        assert torch.equal(layer.forward(x, 0), y1)
    clone.set_seed(6)
    clone.initialize()
    for p, q in zip(layer.parameters(), clone.parameters()):
        assert p.data_ptr() != q.data_ptr()                       # a clone owns its parameters
    layer.clear_cache()
    assert layer.cached_memory_bytes() == 0


def test_microbatch_caches_are_independent():
    layer = dnn.Conv2DLayer(2, 2, 3, 3, 1, 1, 1, 1, False, "c")
    layer.initialize()
    xa, xb = torch.randn(1, 2, 4, 4), torch.randn(1, 2, 4, 4)
    layer.forward(xa, 0)
    layer.forward(xb, 1)
    assert layer.cached_memory_bytes() >= 2 * xa.numel() * 4
    g = torch.ones(1, 2, 4, 4)
    layer.backward(g, 1)
    gb = layer.gradients()[0].clone()
    layer.clear_gradients()
    layer.backward(g, 0)
    ga = layer.gradients()[0].clone()
    single = dnn.Conv2DLayer(2, 2, 3, 3, 1, 1, 1, 1, False, "c")
    single.initialize()
    single.parameters()[0].copy_(layer.parameters()[0])
    single.forward(xb, 0)
    single.backward(g, 0)
    assert torch.allclose(single.gradients()[0], gb, atol=1e-6) and not torch.allclose(ga, gb)


def test_factory_rejects_unknown_types_and_lists_defaults():
    types = LayerFactory.available_types()
    for t in ("dense", "conv2d", "activation", "maxpool2d", "avgpool2d", "dropout", "batchnorm", "groupnorm", "flatten",
              "residual_block"):
        assert t in types
    with pytest.raises(ValueError):
        LayerFactory.create("transformer", {"name": "x", "parameters": {}})


def test_dropout_train_vs_eval_scaling():
    do = dnn.DropoutLayer(0.5, "do")
    do.set_seed(1)
    x = torch.ones(64, 8, 4, 4)
    do.set_training(True)
    y = do.forward(x, 0)
    kept = float((y != 0).float().mean())
    assert 0.4 < kept < 0.6 and torch.allclose(y[y != 0], torch.full_like(y[y != 0], 2.0))
    g = do.backward(torch.ones_like(x), 0)
    assert torch.equal(g != 0, y != 0)                            # backward uses the forward mask
    do.set_training(False)
    assert torch.equal(do.forward(x, 0), x)


# ------------------------------------------------------------------------------------------------ SURVEY appendix A.3
def test_reference_semantics_init_ties_dense_shapes_and_loss_scaling():
    # init: weights AND biases ~ U(-1/sqrt(fan_in), +1/sqrt(fan_in)); BN gamma 1 / beta 0 / mean 0 / var 1
    conv = dnn.Conv2DLayer(16, 32, 3, 3, 1, 1, 1, 1, True, "c")
    conv.set_seed(1)
    conv.initialize()
    bound = 1.0 / (16 * 9) ** 0.5
    w, b = conv.parameters()
    assert float(w.abs().max()) <= bound and float(b.abs().max()) <= bound
    assert float(w.abs().max()) > 0.9 * bound and abs(float(w.mean())) < 0.1 * bound       # fills the interval, centred
    fc = dnn.DenseLayer(100, 10, True, "d")
    fc.set_seed(1)
    fc.initialize()
    assert float(fc.parameters()[0].abs().max()) <= 0.1 and float(fc.parameters()[1].abs().max()) <= 0.1
    bn = dnn.BatchNormLayer(4)
    bn.initialize()
    assert torch.equal(bn.parameters()[0].reshape(-1), torch.ones(4)) and torch.equal(bn.parameters()[1].reshape(-1), torch.zeros(4))
    assert torch.equal(bn.buffers()["running_mean"], torch.zeros(4)) and torch.equal(bn.buffers()["running_var"], torch.ones(4))
    # max-pool ties -> the first maximum in scan order receives the gradient
    mp = dnn.MaxPool2DLayer(2, 2, 2, 2, 0, 0, "mp")
    x = torch.zeros(1, 1, 2, 2)
    mp.forward(x, 0)
    dx = mp.backward(torch.ones(1, 1, 1, 1), 0)
    assert torch.equal(dx, torch.tensor([[[[1.0, 0.0], [0.0, 0.0]]]]))
    # Dense takes any [N, C, H, W] with C*H*W == input_features and emits [N, out, 1, 1]
    y = fc.forward(torch.randn(3, 4, 5, 5), 0)
    assert tuple(y.shape) == (3, 10, 1, 1)
    assert tuple(fc.backward(torch.ones_like(y), 0).shape) == (3, 4, 5, 5)
    # loss = mean over the batch, gradient already scaled by 1/batch
    loss_fn = dnn.SoftmaxCrossEntropyLoss()
    logits = torch.randn(6, 5, 1, 1)
    tgt = torch.nn.functional.one_hot(torch.arange(6) % 5, 5).float().reshape(6, 5, 1, 1)
    loss, g, _ = loss_fn.forward_backward(logits, tgt)
    lr = logits.clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lr.reshape(6, 5), tgt.reshape(6, 5).argmax(1), reduction="mean")
    ref.backward()
    assert abs(float(loss) - float(ref.detach())) < 1e-6 and torch.allclose(g, lr.grad, atol=1e-6)
    # builder auto-names (appendix B.1): every auto-named layer is "<type>_0", residual helper BNs are "bn0"
    m = dnn.SequentialBuilder("n").input([3, 8, 8]).conv2d(4, 3, 3, 1, 1, 1, 1, True).batchnorm().activation("relu") \
        .basic_residual_block(4, 4, 1).build()
    names = [l.name for l in m.get_layers()]
    assert names[0] == "conv2d_0" and names[1] == "batchnorm_0" and names[2] == "activation_0"
    block = m.get_layers()[3]
    assert [l.name for l in block.get_main_path() if l.type() == "batchnorm"] == ["bn0", "bn0"]
