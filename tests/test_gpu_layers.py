"""Layer-level device agnosticity (the reference's layer_device_agnosticity_test.cpp): the same layer object type on
the CPU and on cuda:0 with identical parameters must agree to 1e-3 in FP32-exact mode, including the Conv->MaxPool and
Conv->Dense mini-pipelines and a residual block."""
import os

import pytest
import torch

from dcnn_b200 import nn as dnn
from dcnn_b200.nn.blocks import ResidualBlock

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _exact(monkeypatch):
    monkeypatch.setenv("DCNN_DISABLE_TC", "1")          # FP32-exact CUDA-core GEMMs: tight tolerances apply
    torch.manual_seed(7)
    yield
    torch.cuda.synchronize()


def _pair(make):
    cpu, gpu = make(), make()
    cpu.set_seed(3)
    cpu.initialize()
    gpu.set_seed(3)
    gpu.initialize()
    gpu.set_device("cuda:0")
    for p, q in zip(gpu.parameters(), cpu.parameters()):
        p.copy_(q)
    return cpu, gpu


def _close(a, b, tol=1e-3):
    a, b = a.detach().cpu().float(), b.detach().cpu().float()
    assert a.shape == b.shape
    assert float((a - b).abs().max()) <= tol * max(1.0, float(b.abs().max())), float((a - b).abs().max())


LAYERS = [
    (lambda: dnn.Conv2DLayer(8, 16, 3, 3, 1, 1, 1, 1, True, "c"), (4, 8, 12, 12)),
    (lambda: dnn.Conv2DLayer(3, 8, 5, 5, 2, 2, 2, 2, False, "c"), (2, 3, 17, 17)),
    (lambda: dnn.Conv2DLayer(32, 64, 1, 1, 1, 1, 0, 0, True, "c"), (2, 32, 8, 8)),
    (lambda: dnn.DenseLayer(48, 10, True, "d"), (5, 48, 1, 1)),
    (lambda: dnn.BatchNormLayer(8, 1e-5, 0.1, True, "bn"), (6, 8, 5, 5)),
    (lambda: dnn.GroupNormLayer(2, 8, 1e-5, True, "gn"), (3, 8, 6, 6)),
    (lambda: dnn.MaxPool2DLayer(3, 3, 2, 2, 1, 1, "mp"), (2, 8, 9, 9)),
    (lambda: dnn.AvgPool2DLayer(2, 2, 2, 2, 0, 0, "ap"), (2, 8, 8, 8)),
    (lambda: dnn.ActivationLayer(dnn.create_activation("elu"), "a"), (2, 8, 4, 4)),
    (lambda: dnn.FlattenLayer("f"), (2, 8, 3, 3)),
]


@pytest.mark.parametrize("make,shape", LAYERS, ids=lambda v: v().type() if callable(v) else "x".join(map(str, v)))
def test_layer_same_on_cpu_and_gpu(make, shape):
    cpu, gpu = _pair(make)
    x = torch.randn(*shape)
    yc, yg = cpu.forward(x, 0), gpu.forward(x.cuda(), 0)
    _close(yg, yc)
    g = torch.randn_like(yc)
    dc_, dg = cpu.backward(g, 0), gpu.backward(g.cuda(), 0)
    _close(dg, dc_)
    for a, b in zip(gpu.gradients(), cpu.gradients()):
        _close(a, b)
    for (k, a), (_, b) in zip(gpu.buffers().items(), cpu.buffers().items()):
        _close(a, b)


@pytest.mark.parametrize("tail", ["maxpool", "dense"])
def test_mini_pipelines(tail):
    def make():
        b = dnn.SequentialBuilder("mini").input([4, 10, 10]).conv2d(8, 3, 3, 1, 1, 1, 1, True, "c").activation("relu", "r")
        if tail == "maxpool":
            b.maxpool2d(2, 2, 2, 2, 0, 0, "mp")
        else:
            b.flatten("f").dense(6, True, "fc")
        return b.build()
    cpu, gpu = _pair(make)
    x = torch.randn(3, 4, 10, 10)
    yc, yg = cpu.forward(x, 0), gpu.forward(x.cuda(), 0)
    _close(yg, yc)
    g = torch.randn_like(yc)
    _close(gpu.backward(g.cuda(), 0), cpu.backward(g, 0))
    for a, b in zip(gpu.gradients(), cpu.gradients()):
        _close(a, b)


def test_residual_block_projection():
    def make():
        return ResidualBlock([dnn.Conv2DLayer(8, 16, 3, 3, 2, 2, 1, 1, False, "c1"), dnn.BatchNormLayer(16, 1e-5, 0.1, True, "b1"),
                              dnn.ActivationLayer(dnn.create_activation("relu"), "r"),
                              dnn.Conv2DLayer(16, 16, 3, 3, 1, 1, 1, 1, False, "c2"), dnn.BatchNormLayer(16, 1e-5, 0.1, True, "b2")],
                             [dnn.Conv2DLayer(8, 16, 1, 1, 2, 2, 0, 0, False, "p"), dnn.BatchNormLayer(16, 1e-5, 0.1, True, "pb")],
                             "relu", "rb")
    cpu, gpu = _pair(make)
    x = torch.randn(4, 8, 12, 12)
    yc, yg = cpu.forward(x, 0), gpu.forward(x.cuda(), 0)
    _close(yg, yc)
    g = torch.randn_like(yc)
    _close(gpu.backward(g.cuda(), 0), cpu.backward(g, 0), 2e-3)
    for a, b in zip(gpu.gradients(), cpu.gradients()):
        _close(a, b, 2e-3)


def test_predictor_on_gpu_matches_cpu_eval_and_replays_a_graph():
    """dcnn_b200.serve: folded network + CUDA-graph replay on cuda:0 against the CPU eval-mode network."""
    from dcnn_b200.models import MODEL_ZOO
    from dcnn_b200.serve import Predictor
    cpu = MODEL_ZOO["resnet9_cifar10"]()
    cpu.set_seed(3)
    cpu.initialize()
    cpu.set_training(True)
    for _ in range(2):                                   # move the BatchNorm statistics off their initial values
        cpu.forward(torch.randn(8, 3, 32, 32) * 2 + 0.3)
    cpu.set_training(False)
    gpu = MODEL_ZOO["resnet9_cifar10"]()
    gpu.set_seed(3)
    gpu.initialize()
    gpu.set_device("cuda:0")
    for p, q in zip(gpu.parameters(), cpu.parameters()):
        p.copy_(q)
    for (k, b), (_, c) in zip(sorted(gpu.buffers().items()), sorted(cpu.buffers().items())):
        b.copy_(c)
    p = Predictor(gpu, (3, 32, 32), batch_size=16)
    x = torch.randn(37, 3, 32, 32)
    ref = cpu.forward(x)
    for _ in range(3):                                   # first call warms up and captures, later calls replay
        got = p.predict(x.cuda())
    _close(got.cpu(), ref, 2e-3)
    assert "cuda_graph=True" in p.summary() and "batchnorm 0" in p.summary()
    outs = list(p.stream_batches([x[:16], x[16:32], x[32:]]))
    _close(torch.cat(outs).cpu(), ref, 2e-3)


def _sweep_cases():
    """The CPU sweep tables (tests/test_layer_sweeps_cpu.py) as (constructor, input shape) pairs for the GPU."""
    from tests.test_layer_sweeps_cpu import ACT, BN, CONV, DENSE, GN, POOL
    out = []
    for (N, Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, bias) in CONV:
        out.append((f"conv_{Cin}x{H}x{W}_k{kh}x{kw}_s{sh}{sw}_p{ph}{pw}_oc{Cout}",
                    lambda a=(Cin, Cout, kh, kw, sh, sw, ph, pw, bias): dnn.Conv2DLayer(*a, "c"), (N, Cin, H, W)))
    for (N, fin, fout, bias) in DENSE:
        out.append((f"dense_{fin}_{fout}", lambda a=(fin, fout, bias): dnn.DenseLayer(*a, "d"), (N, fin, 1, 1)))
    for (N, C, H, W, kh, kw, sh, sw, ph, pw) in POOL:
        out.append((f"maxpool_{H}x{W}_k{kh}x{kw}", lambda a=(kh, kw, sh, sw, ph, pw): dnn.MaxPool2DLayer(*a, "p"), (N, C, H, W)))
        out.append((f"avgpool_{H}x{W}_k{kh}x{kw}", lambda a=(kh, kw, sh, sw, ph, pw): dnn.AvgPool2DLayer(*a, "p"), (N, C, H, W)))
    for (N, C, H, W, affine) in BN:
        out.append((f"bn_{C}_{N}x{H}x{W}", lambda a=(C, affine): dnn.BatchNormLayer(a[0], 1e-5, 0.1, a[1], "bn"), (N, C, H, W)))
    for (N, C, G, H, W, affine) in GN:
        out.append((f"gn_{C}_{G}", lambda a=(G, C, affine): dnn.GroupNormLayer(a[0], a[1], 1e-5, a[2], "gn"), (N, C, H, W)))
    for name, _ in ACT:
        out.append((f"act_{name}", lambda a=name: dnn.ActivationLayer(a, "a"), (3, 8, 5, 5)))
    return out


@pytest.mark.parametrize("name,make,shape", _sweep_cases(), ids=[c[0] for c in _sweep_cases()])
def test_layer_sweep_same_on_cpu_and_gpu(name, make, shape):
    """Every case of the CPU sweep, CPU layer vs the same layer on cuda:0 (FP32-exact mode): forward, input gradient,
    parameter gradients, BatchNorm running statistics."""
    cpu, gpu = _pair(make)
    x = torch.randn(*shape) * 1.5 + 0.2
    yc, yg = cpu.forward(x, 0), gpu.forward(x.cuda(), 0)
    _close(yg, yc)
    g = torch.randn_like(yc)
    dc_, dg = cpu.backward(g, 0), gpu.backward(g.cuda(), 0)
    _close(dg, dc_, 2e-3)
    for a, b in zip(gpu.gradients(), cpu.gradients()):
        _close(a, b, 2e-3)
    for (k, a), (_, b) in zip(gpu.buffers().items(), cpu.buffers().items()):
        _close(a, b)


@pytest.mark.parametrize("name", ["mnist_cnn", "all_layers", "resnet9_cifar10", "resnet18_tiny_imagenet"])
def test_gpu_forward_and_training_step_against_the_reference_binary(tmp_path, name):
    """The reference itself as the oracle for the CUDA path: it writes a trained snapshot and its own forward output
    (baseline/_ref/bin/ref_snapshot, CPU FP32), the snapshot is loaded on cuda:0 and must reproduce that output and the
    loss sequence of two SGD steps (FP32-exact mode of this file's fixture)."""
    import subprocess
    from tests.test_reference_interop import BIN, _read_io, _write_io
    if not os.path.exists(BIN):
        pytest.skip("reference subset build absent (bash baseline/build_ref.sh)")
    prefix = str(tmp_path / name)
    r = subprocess.run([BIN, "write", name, prefix, "1"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    x, y_ref, _ = _read_io(prefix + ".io.bin")
    model = dnn.Sequential.from_file(prefix, device="GPU")
    model.set_training(True)
    y = model.forward(x.cuda())
    _close(y, y_ref, 2e-3)
    io_path = str(tmp_path / "x.io")
    _write_io(io_path, x)
    r = subprocess.run([BIN, "train", prefix, io_path, str(tmp_path / "after"), "2", "sgd", "softmax_ce"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    ref_losses = [float(l.split()[1]) for l in r.stdout.splitlines() if l.startswith("loss ")]
    classes = y.shape[1]
    t = torch.zeros(x.shape[0], classes, 1, 1)
    for n in range(x.shape[0]):
        t[n, (3 * n) % classes] = 1.0
    opt = dnn.SGD(0.05, 0.0)
    opt.attach(model.parameters(), model.gradients())
    loss = dnn.SoftmaxCrossEntropyLoss()
    ours = []
    for _ in range(2):
        out = model.forward(x.cuda())
        ours.append(float(loss.compute_loss(out, t.cuda())))
        model.backward(loss.compute_gradient(out, t.cuda()))
        opt.update()
        model.clear_gradients()
    assert abs(ours[0] - ref_losses[0]) < 1e-3 * max(1.0, abs(ref_losses[0]))
    assert abs(ours[1] - ref_losses[1]) < 2e-2 * max(1.0, abs(ref_losses[1])), (ours, ref_losses)
