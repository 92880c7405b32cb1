"""Native CPU math (csrc/host/cpu_math.cpp) against ATen: GEMM in all transpose modes with ragged sizes, dgemm,
elementwise / reduction primitives, im2col/col2im, and the DCNN_CPU_BACKEND=native layer path end to end.
Mirrors the reference's unit_tests/math/sgemm_test.cpp / dgemm_test.cpp and ops tests (tol 1e-4 relative)."""
import pytest
import torch
import torch.nn.functional as F

from dcnn_b200.ops import cpu_native as N


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30))


@pytest.mark.parametrize("M,N_,K", [(1, 1, 1), (5, 7, 3), (6, 16, 8), (97, 33, 129), (128, 128, 128), (200, 17, 300),
                                    (13, 2100, 40), (300, 70, 513)])
@pytest.mark.parametrize("ta,tb", [(False, False), (False, True), (True, False), (True, True)])
def test_sgemm(M, N_, K, ta, tb):
    torch.manual_seed(M * 1000 + N_ * 10 + K)
    a = torch.randn((K, M) if ta else (M, K))
    b = torch.randn((N_, K) if tb else (K, N_))
    want = (a.t() if ta else a) @ (b.t() if tb else b)
    assert _rel(N.gemm(a, b, ta, tb), want) < 1e-5
    c = torch.randn(M, N_)
    want2 = 0.5 * want - 2.0 * c
    N.gemm(a, b, ta, tb, alpha=0.5, beta=-2.0, c=c)
    assert _rel(c, want2) < 1e-5


def test_gemm_beta_zero_ignores_nan_and_dgemm():
    a, b = torch.randn(20, 30), torch.randn(30, 10)
    c = torch.full((20, 10), float("nan"))
    N.gemm(a, b, beta=0.0, c=c)
    assert torch.isfinite(c).all() and _rel(c, a @ b) < 1e-5
    ad, bd = torch.randn(70, 45, dtype=torch.float64), torch.randn(33, 45, dtype=torch.float64)
    assert _rel(N.gemm(ad, bd, trans_b=True), ad @ bd.t()) < 1e-12
    with pytest.raises(ValueError):
        N.gemm(torch.randn(3, 4), torch.randn(5, 6))


def test_elementwise_and_reductions():
    torch.manual_seed(0)
    a, b, c = torch.randn(100003), torch.randn(100003), torch.randn(100003)
    assert torch.equal(N.binary("add", a, b), a + b)
    assert torch.equal(N.binary("sub", a, b), a - b)
    assert torch.equal(N.binary("mul", a, b), a * b)
    assert torch.allclose(N.binary("div", a, b), a / b, rtol=1e-6)
    assert torch.equal(N.binary("max", a, b), torch.maximum(a, b))
    assert torch.equal(N.binary("mul", a, 3.0), a * 3.0)
    assert torch.equal(N.unary("relu", a), torch.relu(a))
    assert torch.equal(N.unary("abs", a), a.abs())
    assert torch.allclose(N.unary("sqrt", a.abs()), a.abs().sqrt(), rtol=1e-6)
    assert torch.allclose(N.unary("exp", a), a.exp(), rtol=1e-5)
    assert torch.allclose(N.fma(a, b, c), a * b + c, atol=1e-6)
    assert torch.allclose(N.fma(a, b, c, "fnmadd"), c - a * b, atol=1e-6)
    y = b.clone()
    N.axpy(0.25, a, y)
    assert torch.allclose(y, b + 0.25 * a, atol=1e-6)
    assert abs(N.dot(a, b) - float(a.double() @ b.double())) < 2e-2
    assert abs(N.sum(a) - float(a.double().sum())) < 1e-3
    m = float(a.mean())
    assert abs(N.sum_squared_diff(a, m) - float(((a.double() - m) ** 2).sum())) < 1e-1
    d = torch.randn(1000, dtype=torch.float64)
    assert abs(N.dot(d, d) - float(d @ d)) < 1e-9
    assert torch.equal(N.fill(torch.empty(77), 1.5), torch.full((77,), 1.5))
    t = torch.randn(70, 45)
    assert torch.equal(N.transpose_2d(t), t.t().contiguous())
    N.set_num_threads(2)
    assert N.get_num_threads() == 2
    assert _rel(N.gemm(torch.ones(64, 64), torch.ones(64, 64)), torch.full((64, 64), 64.0)) < 1e-6


@pytest.mark.parametrize("k,s,p", [(3, 1, 1), (5, 2, 2), (1, 1, 0), (3, 2, 0)])
def test_im2col_col2im(k, s, p):
    x = torch.randn(3, 4, 11, 9)
    col = N.im2col(x, k, k, (s, s), (p, p))
    want = F.unfold(x, k, padding=p, stride=s).permute(1, 0, 2).reshape(col.shape[0], -1)
    assert torch.equal(col, want)
    back = N.col2im(col, tuple(x.shape), k, k, (s, s), (p, p))
    L = col.shape[1] // 3
    want_back = F.fold(col.reshape(col.shape[0], 3, L).permute(1, 0, 2), (11, 9), k, padding=p, stride=s)
    assert torch.allclose(back, want_back, atol=1e-5)


def test_native_backend_layers_match_aten(monkeypatch):
    from dcnn_b200.ops import cpu
    x, w, b = torch.randn(4, 3, 12, 10), torch.randn(8, 3, 3, 3), torch.randn(8, 1, 1, 1)
    dy = torch.randn(4, 8, 6, 5)
    ref_y, _ = cpu.conv2d_forward(x, w, b, (2, 2), (1, 1))
    ref_dx = cpu.conv2d_backward_data(dy, w, x.shape, (2, 2), (1, 1))
    ref_dw = torch.ones(8, 3, 3, 3)
    cpu.conv2d_backward_weight(dy, x, ref_dw, (2, 2), (1, 1))
    x2, w2, b2, dy2 = torch.randn(9, 20), torch.randn(7, 20), torch.randn(7), torch.randn(9, 7)
    ref_d = cpu.dense_forward(x2, w2, b2)
    ref_ddx = cpu.dense_backward_data(dy2, w2)
    ref_ddw = torch.ones(7, 20)
    cpu.dense_backward_weight(dy2, x2, ref_ddw)
    monkeypatch.setenv("DCNN_CPU_BACKEND", "native")
    y, st = cpu.conv2d_forward(x, w, b, (2, 2), (1, 1), want_stats=True)
    assert st is not None and torch.allclose(y, ref_y, atol=1e-5)
    assert torch.allclose(cpu.conv2d_backward_data(dy, w, x.shape, (2, 2), (1, 1)), ref_dx, atol=1e-5)
    dw = torch.ones(8, 3, 3, 3)
    cpu.conv2d_backward_weight(dy, x, dw, (2, 2), (1, 1))
    assert torch.allclose(dw, ref_dw, atol=1e-4)
    assert torch.allclose(cpu.dense_forward(x2, w2, b2), ref_d, atol=1e-5)
    assert torch.allclose(cpu.dense_backward_data(dy2, w2), ref_ddx, atol=1e-5)
    ddw = torch.ones(7, 20)
    cpu.dense_backward_weight(dy2, x2, ddw)
    assert torch.allclose(ddw, ref_ddw, atol=1e-5)


def test_native_backend_trains_mnist_cnn(monkeypatch):
    """The MNIST CNN takes identical steps (to rounding) on the native GEMM path and on ATen."""
    from dcnn_b200 import nn as dnn
    from dcnn_b200.models import create_mnist_trainer
    from dcnn_b200.train import TrainStep
    g = torch.Generator().manual_seed(0)
    x = torch.randn(8, 1, 28, 28, generator=g)
    t = F.one_hot(torch.randint(0, 10, (8,), generator=g), 10).float().reshape(8, 10, 1, 1)

    def run():
        m = create_mnist_trainer()
        m.set_seed(5)
        m.initialize()
        o = dnn.SGD(0.05, 0.9)
        o.attach(m.parameters(), m.gradients())
        step = TrainStep(m, o, dnn.SoftmaxCrossEntropyLoss())
        losses = [float(step(x, t)[0]) for _ in range(3)]
        return losses, [p.clone() for p in m.parameters()]
    l0, p0 = run()
    monkeypatch.setenv("DCNN_CPU_BACKEND", "native")
    l1, p1 = run()
    assert max(abs(a - b) for a, b in zip(l0, l1)) < 1e-4
    for a, b in zip(p0, p1):
        assert torch.allclose(a, b, atol=1e-4)


@pytest.mark.parametrize("relu,res", [(False, False), (True, False), (True, True)])
def test_native_batchnorm_matches_aten(monkeypatch, relu, res):
    from dcnn_b200.ops import cpu
    torch.manual_seed(3)
    x = torch.randn(6, 5, 7, 4) * 2 + 1
    gamma, beta = torch.rand(5, 1, 1, 1) + 0.5, torch.randn(5, 1, 1, 1)
    r = torch.randn_like(x) if res else None
    dout = torch.randn_like(x)

    def run():
        rm, rv = torch.zeros(5), torch.ones(5)
        mean, inv = cpu.bn_finalize(cpu.bn_stats(x), 1e-5, 0.1, rm, rv)
        y = cpu.bn_apply(x, mean, inv, gamma, beta, r, relu)
        dg, db = torch.zeros(5, 1, 1, 1), torch.zeros(5, 1, 1, 1)
        dx, m = cpu.bn_backward(dout, x, y, mean, inv, gamma, dg, db, relu, True)
        return [mean, inv, rm, rv, y, dx, m, dg, db]
    want = run()
    monkeypatch.setenv("DCNN_CPU_BACKEND", "native")
    got = run()
    for a, b in zip(got, want):
        assert torch.allclose(a, b, atol=2e-5, rtol=1e-4), (a - b).abs().max()


@pytest.mark.parametrize("k,s,p", [((2, 2), (2, 2), (0, 0)), ((3, 3), (2, 2), (1, 1)), ((3, 2), (1, 2), (1, 0)),
                                   ((3, 3), (1, 1), (1, 1))])
def test_native_pools_match_aten(monkeypatch, k, s, p):
    from dcnn_b200.ops import cpu
    torch.manual_seed(4)
    x = torch.randn(3, 4, 9, 8)

    def run():
        y, idx = cpu.maxpool_forward(x, k, s, p)
        g = torch.arange(y.numel(), dtype=torch.float32).reshape(y.shape) / 10
        dx = cpu.maxpool_backward(g, idx, x.shape, k, s, p)
        ya = cpu.avgpool_forward(x, k, s, p)
        da = cpu.avgpool_backward(torch.ones_like(ya) * 0.5, x.shape, k, s, p)
        return [y, idx, dx, ya, da]
    want = run()
    monkeypatch.setenv("DCNN_CPU_BACKEND", "native")
    got = run()
    for a, b in zip(got, want):
        assert torch.allclose(a.double(), b.double(), atol=1e-5), (a.double() - b.double()).abs().max()
    # ties: the first maximum in scan order gets the gradient
    y, idx = cpu.maxpool_forward(torch.zeros(1, 1, 2, 2), (2, 2), (2, 2), (0, 0))
    assert int(idx.reshape(-1)[0]) == 0


def test_native_backend_trains_resnet9_like_aten(monkeypatch):
    """Conv + BN + ReLU + residual blocks + pools + Dense through the native kernels == through ATen."""
    from dcnn_b200 import nn as dnn
    from dcnn_b200.models import create_resnet9_cifar10
    from dcnn_b200.train import TrainStep
    g = torch.Generator().manual_seed(0)
    x = torch.randn(4, 3, 32, 32, generator=g)
    t = F.one_hot(torch.randint(0, 10, (4,), generator=g), 10).float().reshape(4, 10, 1, 1)

    def run():
        m = create_resnet9_cifar10()
        m.set_seed(5)
        m.initialize()
        o = dnn.SGD(0.01, 0.9)
        o.attach(m.parameters(), m.gradients())
        step = TrainStep(m, o, dnn.SoftmaxCrossEntropyLoss())
        losses = [float(step(x, t)[0]) for _ in range(2)]
        return losses, [p.clone() for p in m.parameters()], [b.clone() for b in m.buffers().values()]
    l0, p0, b0 = run()
    monkeypatch.setenv("DCNN_CPU_BACKEND", "native")
    l1, p1, b1 = run()
    assert max(abs(a - b) for a, b in zip(l0, l1)) < 1e-3
    for a, b in zip(p0 + b0, p1 + b1):
        assert torch.allclose(a, b, atol=2e-3, rtol=1e-2), (a - b).abs().max()


def test_native_optimizers_and_loss_match_aten(monkeypatch):
    from dcnn_b200.ops import cpu
    torch.manual_seed(5)

    def tensors():
        g = torch.Generator().manual_seed(9)
        p = [torch.randn(33, 7, generator=g), torch.randn(100003, generator=g)]
        gr = [torch.randn(33, 7, generator=g), torch.randn(100003, generator=g)]
        return p, gr, [torch.zeros_like(t) for t in p], [torch.zeros_like(t) for t in p]

    def run():
        out = []
        p, g, v, _ = tensors()
        cpu.sgd_update(p, [t.clone() for t in g], v, 0.1, 0.9)
        cpu.sgd_update(p, g, v, 0.1, 0.9, zero_grads=True)
        out += p + v + g
        p, g, _, _ = tensors()
        cpu.sgd_update(p, g, None, 0.05, 0.0)
        out += p
        for decoupled in (False, True):
            p, g, m, v = tensors()
            cpu.adam_update(p, [t.clone() for t in g], m, v, 1e-2, 0.9, 0.999, 1e-8, 0.1, 0.001, 1e-2, decoupled)
            cpu.adam_update(p, g, m, v, 1e-2, 0.9, 0.999, 1e-8, 0.19, 0.002, 1e-2, decoupled, zero_grads=True)
            out += p + m + v + g
        logits = torch.randn(9, 11, 1, 1, generator=torch.Generator().manual_seed(21))
        tgt = F.one_hot(torch.arange(9) % 11, 11).float().reshape(9, 11, 1, 1)
        loss, grad, correct = cpu.loss_forward_backward("softmax_crossentropy", logits, tgt)
        loss2, none, _ = cpu.loss_forward_backward("logsoftmax_crossentropy", logits, tgt, want_grad=False)
        assert none is None
        out += [loss.reshape(1), grad, correct.reshape(1).float(), loss2.reshape(1)]
        return out
    want = run()
    monkeypatch.setenv("DCNN_CPU_BACKEND", "native")
    got = run()
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert torch.allclose(a.float(), b.float(), atol=1e-6, rtol=1e-5), (a.float() - b.float()).abs().max()


@pytest.mark.parametrize("M,N_,K", [(8, 25, 36864), (1, 1, 4096), (48, 80, 5000), (64, 64, 4097)])
def test_sgemm_long_k_nt_path(M, N_, K):
    """Small M x N with a long K and both operands K-contiguous (a conv weight gradient) takes the split-K dot path."""
    torch.manual_seed(K)
    a, b = torch.randn(M, K), torch.randn(N_, K)
    want = (a.double() @ b.double().t()).float()
    got = N.gemm(a, b, trans_b=True)
    assert _rel(got, want) < 1e-5
    c = torch.randn(M, N_)
    want2 = (0.5 * want.double() + 2.0 * c.double()).float()
    N.gemm(a, b, trans_b=True, alpha=0.5, beta=2.0, c=c)
    assert _rel(c, want2) < 1e-5
    nanc = torch.full((M, N_), float("nan"))
    N.gemm(a, b, trans_b=True, beta=0.0, c=nanc)
    assert torch.isfinite(nanc).all()


@pytest.mark.parametrize("shape", [
    # N, C, H, W, OC, k, stride, pad
    (64, 1, 28, 28, 8, 5, 1, 0),      # one image per block (OH*OW = 576): GEMM writes NCHW in place
    (64, 16, 8, 8, 48, 5, 1, 0),      # 64 images per block, AVX-512 tile path
    (37, 3, 9, 11, 7, 3, 2, 1),       # strided, padded, ragged last block
    (70, 2, 40, 40, 4, 3, 1, 1),      # two images per block, odd block count
    (3, 4, 6, 6, 5, 3, 1, 2),         # padding wider than half the kernel
    (2, 2, 5, 5, 3, 5, 1, 2),         # kernel as large as the image
    (9, 3, 7, 7, 4, 3, 3, 0),         # stride == kernel
    (2, 1, 1, 1, 2, 5, 2, 2),         # 1x1 input living entirely on padding taps
])
@pytest.mark.parametrize("threads", [1, 4])
def test_blocked_native_convolution_matches_aten(shape, threads):
    """csrc/host/cpu_conv.cpp: forward, data gradient and (accumulating) weight gradient against ATen."""
    B, C, H, W, OC, k, st, pd = shape
    old = N.get_num_threads()
    N.set_num_threads(threads)
    try:
        g = torch.Generator().manual_seed(B * 131 + k)
        x, w, b = torch.randn(B, C, H, W, generator=g), torch.randn(OC, C, k, k, generator=g), torch.randn(OC, generator=g)
        y = N.conv2d_forward(x, w, b, (st, st), (pd, pd))
        y_ref = F.conv2d(x, w, b, st, pd)
        assert y.shape == y_ref.shape and torch.allclose(y, y_ref, atol=2e-4, rtol=1e-4)
        assert torch.allclose(N.conv2d_forward(x, w, None, (st, st), (pd, pd)), F.conv2d(x, w, None, st, pd),
                              atol=2e-4, rtol=1e-4)
        dy = torch.randn(y_ref.shape, generator=g)
        dx = N.conv2d_backward_data(dy, w, x.shape, (st, st), (pd, pd))
        assert torch.allclose(dx, torch.nn.grad.conv2d_input(x.shape, w, dy, st, pd), atol=3e-4, rtol=1e-4)
        dw = torch.full_like(w, 0.5)
        N.conv2d_backward_weight(dy, x, dw, (st, st), (pd, pd))          # accumulates
        dw_ref = 0.5 + torch.nn.grad.conv2d_weight(x, w.shape, dy, st, pd)
        assert torch.allclose(dw, dw_ref, atol=1e-3 * float(dw_ref.abs().max()), rtol=1e-4)
        dwt = torch.zeros(C, OC, k, k).permute(1, 0, 2, 3)                       # non-contiguous gradient tensor
        N.conv2d_backward_weight(dy, x, dwt, (st, st), (pd, pd))
        assert torch.allclose(dwt, dw_ref - 0.5, atol=1e-3 * float(dw_ref.abs().max()), rtol=1e-4)
    finally:
        N.set_num_threads(old)
