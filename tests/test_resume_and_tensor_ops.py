import torch
import torch.nn.functional as F

from dcnn_b200 import nn as dnn
from dcnn_b200.io import load_training_state, save_training_state
from dcnn_b200.models import create_mnist_trainer
from dcnn_b200.ops import tensor_ops as T
from dcnn_b200.train import TrainStep


def _data():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(8, 1, 28, 28, generator=g)
    t = F.one_hot(torch.randint(0, 10, (8,), generator=g), 10).float().reshape(8, 10, 1, 1)
    return x, t


def test_resume_continues_bit_identically(tmp_path):
    x, t = _data()

    def fresh():
        m = create_mnist_trainer()
        m.set_seed(1)
        m.initialize()
        o = dnn.Adam(1e-2)
        o.attach(m.parameters(), m.gradients())
        return m, o, TrainStep(m, o, dnn.SoftmaxCrossEntropyLoss())
    m, o, step = fresh()
    for _ in range(3):
        step(x, t)
    path = str(tmp_path / "ck" / "mnist")
    save_training_state(path, m, o, epoch=2, extra={"note": "hi"})
    for _ in range(2):
        step(x, t)
    want = [p.clone() for p in m.parameters()] + list(m.buffers().values())
    m2, o2, step2 = fresh()
    blob = load_training_state(path, m2, o2)
    assert blob["epoch"] == 2 and blob["extra"]["note"] == "hi" and o2.t == 3
    for _ in range(2):
        step2(x, t)
    got = [p for p in m2.parameters()] + list(m2.buffers().values())
    for a, b in zip(got, want):
        assert torch.equal(a, b)


def test_tensor_ops_match_reference_semantics():
    x = torch.randn(2, 3, 6, 5)
    cols = T.im2col(x, 3, 3, 1, 1, 1, 1)
    assert cols.shape == (27, 2 * 6 * 5)
    w = torch.randn(4, 3, 3, 3)
    y = (w.reshape(4, 27) @ cols).reshape(4, 2, 6, 5).permute(1, 0, 2, 3)
    assert torch.allclose(y, F.conv2d(x, w, padding=1), atol=1e-5)
    back = T.col2im(cols, 2, 3, 6, 5, 3, 3, 1, 1, 1, 1)
    ones = T.col2im(torch.ones_like(cols), 2, 3, 6, 5, 3, 3, 1, 1, 1, 1)
    assert torch.allclose(back / ones, x, atol=1e-5)
    p = T.pad(x, 2, 1)
    assert p.shape == (2, 3, 10, 7) and torch.equal(T.unpad(p, 2, 1), x)
    assert T.crop(x, 1, 1, 3, 2).shape == (2, 3, 3, 2)
    assert T.slice_channels(x, 1, 2).shape == (2, 2, 6, 5) and T.slice_batch(x, 1, 2).shape == (1, 3, 6, 5)
    parts = T.split(torch.arange(10).reshape(10, 1, 1, 1), 3)
    assert [q.shape[0] for q in parts] == [3, 3, 4]
    s = T.apply_softmax(x.clone())
    assert torch.allclose(s.sum(1), torch.ones(2, 6, 5), atol=1e-6)
    a, b = torch.randn(4, 6), torch.randn(6, 5)
    assert torch.allclose(T.gemm(a, b), a @ b, atol=1e-6)
    assert torch.allclose(T.gemm(a, b.t().contiguous(), trans_b=True, alpha=2.0), 2 * a @ b, atol=1e-5)
    c = torch.ones(4, 5)
    T.gemm(a, b, beta=0.5, c=c)
    assert torch.allclose(c, a @ b + 0.5, atol=1e-5)
    yv = torch.ones(3)
    T.axpy(2.0, torch.tensor([1.0, 2.0, 3.0]), yv)
    assert torch.equal(yv, torch.tensor([3.0, 5.0, 7.0]))
    assert float(T.dot_product(torch.ones(4), torch.arange(4.0))) == 6.0
    assert torch.equal(T.cnhw_to_nchw(T.nchw_to_cnhw(x)), x)


def test_packed_tap_stem_formulation():
    """Index convention of the packed-tap stem path (csrc/conv_simt.cu pack_taps_kernel, ops/cuda.py _packed_weights):
    P[n,oh,ow,(r*S+s)*Cin+c] against the O-H-W-I weight flattening reproduces conv2d, and P^T dY reproduces dW."""
    N, Cin, H, W, Cout, R, S, st, pd = 2, 3, 9, 8, 8, 3, 3, 2, 1
    x = torch.randn(N, Cin, H, W)
    w = torch.randn(Cout, Cin, R, S)
    OH, OW = (H + 2 * pd - R) // st + 1, (W + 2 * pd - S) // st + 1
    Kp = 32
    P = torch.zeros(N, OH, OW, Kp)
    for n in range(N):
        for oh in range(OH):
            for ow in range(OW):
                for r in range(R):
                    for s in range(S):
                        ih, iw = oh * st - pd + r, ow * st - pd + s
                        if 0 <= ih < H and 0 <= iw < W:
                            P[n, oh, ow, (r * S + s) * Cin:(r * S + s + 1) * Cin] = x[n, :, ih, iw]
    wp = torch.zeros(Cout, Kp)
    wp[:, :R * S * Cin] = w.permute(0, 2, 3, 1).reshape(Cout, -1)
    y = (P.reshape(-1, Kp) @ wp.t()).reshape(N, OH, OW, Cout).permute(0, 3, 1, 2)
    assert torch.allclose(y, F.conv2d(x, w, stride=st, padding=pd), atol=1e-5)
    dy = torch.randn(N, Cout, OH, OW)
    dwp = dy.permute(0, 2, 3, 1).reshape(-1, Cout).t() @ P.reshape(-1, Kp)
    dw = dwp[:, :R * S * Cin].reshape(Cout, R, S, Cin).permute(0, 3, 1, 2)
    want = torch.nn.grad.conv2d_weight(x, w.shape, dy, stride=st, padding=pd)
    assert torch.allclose(dw, want, atol=1e-4)


def test_training_loop_resumes_bit_identically(tmp_path, monkeypatch):
    """3 epochs in one go == 2 epochs (CHECKPOINT_EVERY=1) + a fresh process state resumed for the third: same weights,
    same statistics, same history -- shuffling and augmentation draws included."""
    from dcnn_b200.data import AugmentationBuilder, SyntheticDataLoader
    from dcnn_b200.train import TrainingConfig, train_classification_model

    def setup(epochs, snap):
        cfg = TrainingConfig()
        cfg.epochs, cfg.batch_size, cfg.snapshot_dir, cfg.progress_print_interval = epochs, 8, str(snap), 1000
        cfg.lr_decay_interval, cfg.lr_decay_factor = 1, 0.5
        tr, te = SyntheticDataLoader(32, (1, 28, 28), 10, seed=0), SyntheticDataLoader(16, (1, 28, 28), 10, seed=1)
        tr.set_seed(5)
        te.set_seed(6)
        tr.set_augmentation(AugmentationBuilder().horizontal_flip(0.5).brightness(0.5, 0.2).build())
        tr.augmentation.set_seed(7)
        m = create_mnist_trainer()
        m.set_seed(1)
        m.initialize()
        return cfg, tr, te, m, dnn.Adam(1e-2)
    torch.manual_seed(0)
    cfg, tr, te, m_full, o = setup(3, tmp_path / "a")
    h_full = train_classification_model(m_full, tr, te, o, dnn.SoftmaxCrossEntropyLoss(), cfg)
    monkeypatch.setenv("CHECKPOINT_EVERY", "1")
    torch.manual_seed(0)
    cfg, tr, te, m1, o = setup(2, tmp_path / "b")
    h1 = train_classification_model(m1, tr, te, o, dnn.SoftmaxCrossEntropyLoss(), cfg)
    ck = str(tmp_path / "b" / (m1.name() + ".checkpoint"))
    monkeypatch.delenv("CHECKPOINT_EVERY")
    monkeypatch.setenv("RESUME_FROM", ck)
    cfg, tr, te, m2, o = setup(3, tmp_path / "c")          # brand-new objects, as after a restart
    h2 = train_classification_model(m2, tr, te, o, dnn.SoftmaxCrossEntropyLoss(), cfg)
    assert [h["epoch"] for h in h1] == [1, 2] and [h["epoch"] for h in h2] == [3]
    assert abs(h2[0]["train_loss"] - h_full[2]["train_loss"]) < 1e-7 and h2[0]["val_acc"] == h_full[2]["val_acc"]
    for a, b in zip(list(m2.parameters()) + list(m2.buffers().values()),
                    list(m_full.parameters()) + list(m_full.buffers().values())):
        assert torch.equal(a, b)
