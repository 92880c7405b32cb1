"""Inference path: BatchNorm folding equals the eval-mode network (plain stacks, residual blocks with identity and
projection shortcuts, conv biases, non-affine BN), and the Predictor pads / splits requests correctly."""
import pytest
import torch

from dcnn_b200 import nn
from dcnn_b200.models import MODEL_ZOO
from dcnn_b200.nn.layers import BatchNormLayer, Conv2DLayer
from dcnn_b200.serve import Predictor, count_layers, fold_batchnorm


def _trained(make, shape, steps=3):
    """A model whose BatchNorm statistics and affine parameters are no longer at their initial values."""
    torch.manual_seed(0)
    m = make()
    m.set_seed(3)
    m.initialize()
    opt = nn.SGD(0.05, 0.9)
    opt.attach(m.parameters(), m.gradients())
    for _ in range(steps):
        x = torch.randn(8, *shape) * 2 + 0.5
        y = m.forward(x)
        m.backward(torch.randn_like(y))
        opt.update(zero_grads=True)
    m.set_training(False)
    return m


def _custom():
    b = nn.SequentialBuilder("foldme").input([3, 12, 12])
    b.conv2d(6, 3, 3, 1, 1, 1, 1, True, "c_bias").batchnorm(1e-3, 0.1, True, "bn_a").activation("relu", "r")
    b.conv2d(6, 1, 1, 1, 1, 0, 0, False, "c_nobias").batchnorm(1e-5, 0.1, False, "bn_noaffine")
    b.basic_residual_block(6, 6, 1, "rb_id").basic_residual_block(6, 12, 2, "rb_proj")
    b.conv2d(12, 3, 3, 1, 1, 1, 1, True, "c_tail").activation("tanh", "t")      # a conv without BN stays as it is
    b.batchnorm(1e-5, 0.1, True, "bn_alone")                                      # a BN without conv stays as it is
    b.avgpool2d(6, 6).flatten().dense(4)
    return b.build()


@pytest.mark.parametrize("make,shape", [(_custom, (3, 12, 12)), (MODEL_ZOO["resnet9_cifar10"], (3, 32, 32)),
                                        (MODEL_ZOO["mnist_cnn"], (1, 28, 28))])
def test_folded_network_equals_eval_network(make, shape):
    m = _trained(make, shape)
    f = fold_batchnorm(m)
    assert not f.is_training()
    x = torch.randn(5, *shape)
    y, yf = m.forward(x), f.forward(x)
    assert yf.shape == y.shape and float((y - yf).abs().max()) < 2e-4 * max(1.0, float(y.abs().max()))
    convs, bns = count_layers(m, Conv2DLayer), count_layers(m, BatchNormLayer)
    assert count_layers(f, Conv2DLayer) == convs
    if make is _custom:
        assert bns == 8 and count_layers(f, BatchNormLayer) == 1          # only the stand-alone BN survives
    else:
        assert count_layers(f, BatchNormLayer) == 0 < bns
    # the source model is untouched and still trainable
    m.set_training(True)
    m.backward(torch.randn_like(m.forward(x)))


def test_fold_needs_an_initialised_model():
    with pytest.raises(RuntimeError):
        fold_batchnorm(MODEL_ZOO["mnist_cnn"]())


def test_predictor_pads_splits_and_reduces():
    m = _trained(MODEL_ZOO["mnist_cnn"], (1, 28, 28))
    p = Predictor(m, (1, 28, 28), batch_size=8)
    assert "batchnorm 0" in p.summary()
    x = torch.randn(19, 1, 28, 28)
    ref = m.forward(x)
    got = p.predict(x)                                    # 8 + 8 + 3 (padded)
    assert got.shape == ref.shape and torch.allclose(got, ref, atol=2e-4, rtol=1e-3)
    assert torch.allclose(p.predict(x[:3]), ref[:3], atol=2e-4, rtol=1e-3)
    assert p.predict(x[:0]).shape[0] == 0
    assert torch.equal(p.predict_classes(x), ref.reshape(19, -1).argmax(1))
    pr = p.predict_proba(x)
    assert pr.shape == (19, 10) and torch.allclose(pr.sum(1), torch.ones(19), atol=1e-5)
    outs = list(p.stream_batches([x[:8], x[8:16], x[16:]]))
    assert [o.shape[0] for o in outs] == [8, 8, 3] and torch.allclose(torch.cat(outs), ref, atol=2e-4, rtol=1e-3)
    with pytest.raises(ValueError):
        p.predict(torch.randn(2, 3, 28, 28))
    unfolded = Predictor(m, (1, 28, 28), batch_size=4, fold_bn=False)
    assert torch.allclose(unfolded.predict(x), ref, atol=1e-5) and "batchnorm 3" in unfolded.summary()
