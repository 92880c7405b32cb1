"""ResidualBlock and Sequential assembly tests on CPU (the reference's residual_block_test.cpp and
sequential_residual_block_test.cpp kinds): identity / projection / strided shortcuts against autograd, output-shape
inference, JSON config round-trip incl. nested blocks, ResNet-like stacks, gradient finiteness, container operations,
split/partition behaviour and training-mode propagation."""
import json

import pytest
import torch
import torch.nn.functional as F

import dcnn_b200 as dc
from dcnn_b200 import nn as dnn
from dcnn_b200.nn.blocks import ResidualBlock, residual_block
from dcnn_b200.nn.factory import LayerFactory
from dcnn_b200.partitioner import NaivePartitioner

torch.manual_seed(0)


def _conv(ci, co, k, s, p, name):
    return dnn.Conv2DLayer(ci, co, k, k, s, s, p, p, False, name)


def _torch_block(x, params, stride, projection, act):
    """params: main conv1, conv2 weights (+ shortcut conv weight)."""
    h = F.conv2d(x, params[0], stride=stride, padding=1)
    h = F.relu(h)
    h = F.conv2d(h, params[1], padding=1)
    sc = F.conv2d(x, params[2], stride=stride) if projection else x
    out = h + sc
    return {"relu": F.relu, "linear": lambda t: t, "tanh": torch.tanh}[act](out)


@pytest.mark.parametrize("stride,ci,co,act", [(1, 4, 4, "relu"), (2, 4, 8, "relu"), (1, 4, 6, "linear"), (1, 3, 3, "tanh")])
def test_residual_block_shortcuts_against_autograd(stride, ci, co, act):
    projection = stride != 1 or ci != co
    main = [_conv(ci, co, 3, stride, 1, "c1"), dnn.ActivationLayer(dnn.create_activation("relu"), "r"),
            _conv(co, co, 3, 1, 1, "c2")]
    sc = [_conv(ci, co, 1, stride, 0, "sc")] if projection else []
    blk = ResidualBlock(main, sc, act, "blk")
    blk.set_seed(3)
    blk.initialize()
    x = torch.randn(2, ci, 8, 8)
    y = blk.forward(x, 0)
    assert list(y.shape) == blk.compute_output_shape([2, ci, 8, 8]) == [2, co, 8 // stride, 8 // stride]
    g = torch.randn_like(y)
    dx = blk.backward(g, 0)
    xr = x.clone().requires_grad_(True)
    ps = [p.clone().requires_grad_(True) for p in blk.parameters()]
    yr = _torch_block(xr, ps, stride, projection, act)
    assert torch.allclose(y, yr, atol=1e-5)
    grads = torch.autograd.grad(yr, [xr] + ps, g)
    assert torch.allclose(dx, grads[0], atol=1e-4)
    for mine, ref in zip(blk.gradients(), grads[1:]):
        assert torch.allclose(mine, ref, atol=1e-4)
    assert len(blk.get_main_path()) == 3 and len(blk.get_shortcut_path()) == len(sc)


def test_residual_block_config_roundtrip_nested_json():
    blk = residual_block([_conv(4, 8, 3, 2, 1, "c1"), dnn.BatchNormLayer(8, 1e-5, 0.1, True, "bn")],
                         [_conv(4, 8, 1, 2, 0, "proj")], "relu", "rb")
    cfg = blk.get_config()
    text = json.dumps(cfg)
    again = LayerFactory.create("residual_block", json.loads(text))
    assert again.get_config() == cfg
    assert [l.type() for l in again.get_main_path()] == ["conv2d", "batchnorm"]
    assert again.get_shortcut_path()[0].get_config() == blk.get_shortcut_path()[0].get_config()
    for b in (blk, again):
        b.set_seed(9)
        b.initialize()
        b.set_training(False)
    x = torch.randn(1, 4, 6, 6)
    assert torch.equal(blk.forward(x, 0), again.forward(x, 0))


def test_identity_block_requires_matching_shapes():
    blk = ResidualBlock([_conv(4, 6, 3, 1, 1, "c")], [], "relu", "bad")
    blk.initialize()
    with pytest.raises(Exception):
        blk.forward(torch.randn(1, 4, 5, 5), 0)


def _resnet_like(blocks):
    b = dnn.SequentialBuilder("mini_resnet").input([3, 16, 16]).conv2d(8, 3, 3, 1, 1, 1, 1, False, "stem") \
        .batchnorm(1e-5, 0.1, True, "bn").activation("relu", "relu")
    ch = 8
    for i, (kind, out, stride) in enumerate(blocks):
        if kind == "basic":
            b.basic_residual_block(ch, out, stride, f"b{i}")
        else:
            b.bottleneck_residual_block(ch, out // 4, out, stride, f"b{i}")
        ch = out
    return b.avgpool2d(16 // (2 ** sum(1 for _, _, s in blocks if s == 2)),
                       16 // (2 ** sum(1 for _, _, s in blocks if s == 2)), 1, 1, 0, 0, "gap").flatten("f") \
        .dense(5, True, "fc").build()


@pytest.mark.parametrize("blocks", [
    [("basic", 8, 1), ("basic", 16, 2)],
    [("bottleneck", 32, 1), ("bottleneck", 32, 1), ("bottleneck", 64, 2)],
    [("basic", 8, 1), ("bottleneck", 32, 2), ("basic", 32, 1)],
])
def test_resnet_like_stacks_shapes_gradients_and_json(blocks):
    m = _resnet_like(blocks)
    m.set_seed(4)
    m.initialize()
    assert m.compute_output_shape([2, 3, 16, 16])[:2] == [2, 5]
    x = torch.randn(2, 3, 16, 16)
    # JSON round trip first (fresh running statistics on both sides): same seed + config -> same function
    cfg = m.get_config()
    twin = dc.Sequential.load_from_config(json.loads(json.dumps(cfg)))
    assert twin.get_config() == cfg and twin.layer_size() == m.layer_size()
    twin.set_seed(4)
    twin.initialize()
    m.set_training(False)
    twin.set_training(False)
    assert torch.equal(m.forward(x, 0), twin.forward(x, 0))
    m.set_training(True)
    y = m.forward(x, 0)
    assert y.shape[:2] == (2, 5) and torch.isfinite(y).all()
    dx = m.backward(torch.randn_like(y), 0)
    assert dx.shape == x.shape and torch.isfinite(dx).all()
    for g in m.gradients():
        assert torch.isfinite(g).all()
    norms = [float(g.norm()) for g in m.gradients() if g.numel() > 1]
    assert max(norms) > 0 and max(norms) < 1e4


def test_sequential_container_operations():
    m = dc.Sequential("s")
    a, b, c = dnn.FlattenLayer("a"), dnn.DropoutLayer(0.1, "b"), dnn.FlattenLayer("c")
    m.add(a)
    m.add(c)
    m.insert(1, b)
    assert [l.name for l in m.get_layers()] == ["a", "b", "c"] and m.layer_size() == len(m) == 3
    m.remove(1)
    assert [l.name for l in m.get_layers()] == ["a", "c"]
    m.set_name("renamed")
    assert m.name() == "renamed"
    m.clear()
    assert m.layer_size() == 0
    with pytest.raises(RuntimeError):
        m.forward(torch.randn(1, 1, 1, 1))
    with pytest.raises(RuntimeError):
        m.backward(torch.randn(1, 1, 1, 1))


def test_training_mode_propagates_into_blocks():
    m = _resnet_like([("basic", 8, 1)])
    m.initialize()
    m.set_training(False)
    inner = [l for l in m.get_layers() if isinstance(l, ResidualBlock)][0]
    assert not m.is_training() and all(not l.is_training() for l in inner.get_main_path())
    x = torch.randn(4, 3, 16, 16)
    before = {k: v.clone() for k, v in m.buffers().items()}
    m.forward(x, 0)
    for k, v in m.buffers().items():
        assert torch.equal(v, before[k])                  # eval mode leaves running statistics alone
    m.set_training(True)
    m.forward(x, 0)
    assert any(not torch.equal(v, before[k]) for k, v in m.buffers().items())


def test_split_matches_partitioner_and_reassembles():
    m = _resnet_like([("basic", 8, 1), ("basic", 16, 2)])
    m.set_seed(1)
    m.initialize()
    parts = NaivePartitioner().get_partitions(m.get_layers(), 3)
    assert parts[0].start_layer == 0 and parts[-1].end_layer == m.layer_size()
    assert all(p.end_layer == q.start_layer for p, q in zip(parts, parts[1:]))
    stages = m.split(parts)
    assert sum(s.layer_size() for s in stages) == m.layer_size()
    x = torch.randn(2, 3, 16, 16)
    m.set_training(False)
    want = m.forward(x, 0)
    h = x
    for s, p in zip(stages, parts):
        s.initialize()
        for dst, src in zip(s.parameters(), m.parameters(p)):
            dst.copy_(src)
        s.set_training(False)
        h = s.forward(h, 0)
    assert torch.allclose(h, want, atol=1e-6)             # the stages chained == the whole model
    # parameters(partition) enumerates exactly the stage's own parameters, in order
    assert sum(len(m.parameters(p)) for p in parts) == len(m.parameters())
