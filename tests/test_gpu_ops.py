"""GPU kernel tests: every CUDA op against the CPU FP32 oracle of the same op (the reference's strategy:
unit_tests/cuda_*_ops_test.cpp compare GPU results with the CPU implementation, tol 1e-4; layer-level 1e-3).
Tensor-core paths use TF32 inputs with FP32 accumulation, so they get a relative bound instead (documented in
DESIGN.md); DCNN_DISABLE_TC=1 gives the FP32-exact CUDA-core path, tested with the tight bound."""
import os

import pytest
import torch
import torch.nn.functional as F

from dcnn_b200 import ops
from dcnn_b200.ops import cpu as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def cl(t):
    return t.to(DEV).contiguous(memory_format=torch.channels_last)


def rel_err(a, b):
    a, b = a.detach().cpu().float(), b.detach().cpu().float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def check_close(a, b, tol, what=""):
    """Global max-norm bound, and for NCHW tensors the same bound PER CHANNEL (a wrong channel of small magnitude must
    not hide behind the largest one) plus a relative L2 bound."""
    e = rel_err(a, b)
    assert e < tol, f"{what}: rel err {e:.3e} >= {tol}"
    a, b = a.detach().cpu().float(), b.detach().cpu().float()
    if b.dim() == 4 and b.shape[1] > 1 and a.shape == b.shape:
        floor = 0.05 * float(b.abs().max()) + 1e-12          # channels that are ~0 everywhere are judged on the floor
        per = (a - b).abs().amax((0, 2, 3)) / b.abs().amax((0, 2, 3)).clamp_min(floor)
        c = int(per.argmax())
        assert float(per[c]) < 2 * tol, f"{what}: channel {c} rel err {float(per[c]):.3e} >= {2 * tol}"
    l2 = float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
    assert l2 < tol, f"{what}: rel L2 {l2:.3e} >= {tol}"


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(1234)
    yield
    torch.cuda.synchronize()


# ----------------------------------------------------------------------------------------- TMA semantics probes
def _probe(x_nhwc, bn, bh, bw, mul, c0, c1, c2, c3):
    from dcnn_b200._lib import I, P, check, lib
    N, H, W, C = x_nhwc.shape
    out = torch.empty(bn * bh * bw * 32, device=DEV)
    check(lib().dcnn_tma_probe(P(x_nhwc.data_ptr()), I(N), I(H), I(W), I(C), I(bn), I(bh), I(bw), I(mul), I(mul),
                               I(c0), I(c1), I(c2), I(c3), P(out.data_ptr()), P(0)), "tma_probe")
    torch.cuda.synchronize()
    raw = out.cpu().reshape(bn * bh * bw, 8, 4)  # rows x 16-byte chunks x 4 floats
    rows = torch.arange(raw.shape[0])
    unsw = torch.empty_like(raw)
    for j in range(8):  # undo the 128B swizzle: physical chunk = logical chunk XOR (row % 8)
        unsw[:, j] = raw[rows, (j ^ (rows % 8))]
    return unsw.reshape(bn, bh, bw, 32)


def test_tma_negative_coords_zero_fill():
    x = torch.randn(2, 6, 6, 32, device=DEV)
    got = _probe(x, 1, 8, 16, 1, 0, -1, -1, 1)
    exp = torch.zeros(1, 8, 16, 32)
    xc = x.cpu()
    for h in range(8):
        for w in range(16):
            ih, iw = h - 1, w - 1
            if 0 <= ih < 6 and 0 <= iw < 6:
                exp[0, h, w] = xc[1, ih, iw]
    assert torch.equal(got, exp)


def test_tma_element_strides():
    x = torch.randn(2, 16, 16, 32, device=DEV)
    got = _probe(x, 2, 8, 8, 2, 0, -1, -1, 0)
    exp = torch.zeros(2, 8, 8, 32)
    xc = x.cpu()
    for n in range(2):
        for h in range(8):
            for w in range(8):
                ih, iw = 2 * h - 1, 2 * w - 1
                if 0 <= ih < 16 and 0 <= iw < 16:
                    exp[n, h, w] = xc[n, ih, iw]
    assert torch.equal(got, exp)


# ----------------------------------------------------------------------------------------- convolution
CONV_CASES = [
    # N, C, H, W, OC, k, s, p, bias
    (4, 64, 16, 16, 128, 3, 1, 1, True),
    (2, 32, 8, 8, 64, 1, 1, 0, False),
    (8, 64, 32, 32, 64, 3, 1, 1, False),
    (3, 128, 14, 14, 256, 3, 1, 1, True),
    (5, 256, 7, 7, 96, 1, 1, 0, True),
    (2, 64, 4, 4, 512, 3, 1, 1, False),
    (4, 64, 16, 16, 128, 3, 2, 1, False),
    (4, 128, 16, 16, 256, 1, 2, 0, False),
    (2, 32, 30, 30, 64, 7, 2, 3, True),
    (2, 3, 32, 32, 64, 3, 1, 1, True),       # RGB stem: packed taps (K = 27 -> 32) on the tensor cores
    (2, 3, 40, 40, 64, 7, 2, 3, True),       # ImageNet stem: packed taps (K = 147 -> 160), strided
    (2, 1, 28, 28, 8, 5, 1, 0, True),        # MNIST stem: K = 25 -> 32
    (2, 8, 12, 12, 16, 5, 1, 0, True),       # channel counts that are multiples of 4 but not of 32:
    (2, 16, 9, 9, 48, 5, 1, 0, True),        #   the TMA unit pads the reduction dimension with zeros
    (2, 100, 10, 10, 200, 3, 1, 1, True),
    (3, 20, 9, 9, 12, 1, 1, 0, False),
    (2, 36, 12, 12, 40, 3, 2, 1, False),
    (2, 6, 8, 8, 10, 3, 1, 1, True),         # neither multiple of 4: CUDA-core kernels
    (3, 32, 13, 13, 32, 3, 1, 1, False),     # wgrad with two taps per CTA (Cout <= 64, stride 1, same-size output):
    (2, 40, 10, 10, 48, 5, 1, 2, False),     #   odd image sizes, 25 taps (an unpaired last tap), channels % 32 != 0,
    (2, 256, 8, 8, 64, 3, 1, 1, False),      #   several input-channel tiles
]


def _tol(C):
    # every shape the tensor cores can take (TF32 inputs) gets the TF32 bound; the CUDA-core cases are far inside it
    return 4e-3 if os.environ.get("DCNN_DISABLE_TC", "0") != "1" else 2e-5


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "x".join(map(str, c)))
def test_conv_forward(case):
    N, C, H, W, OC, k, s, p, bias = case
    x = torch.randn(N, C, H, W)
    w = torch.randn(OC, C, k, k) / (C * k * k) ** 0.5
    b = torch.randn(OC, 1, 1, 1) if bias else None
    y_ref, st_ref = ref.conv2d_forward(x, w, b, (s, s), (p, p), True)
    y, st = ops.conv2d_forward(cl(x), cl(w), None if b is None else b.to(DEV), (s, s), (p, p), True)
    check_close(y, y_ref, _tol(C), "conv fprop")
    rm, rv = torch.zeros(OC, device=DEV), torch.ones(OC, device=DEV)
    mean, inv = ops.bn_finalize(st, 1e-5, 0.1, rm, rv)
    mean_ref, inv_ref = ref.bn_finalize(st_ref, 1e-5, 0.1, torch.zeros(OC), torch.ones(OC))
    assert (mean.cpu() - mean_ref).abs().max() < 5e-3
    check_close(inv[0], inv_ref, 1e-2, "epilogue BN invstd")


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "x".join(map(str, c)))
def test_conv_backward_data(case):
    N, C, H, W, OC, k, s, p, _ = case
    w = torch.randn(OC, C, k, k) / (C * k * k) ** 0.5
    oh = (H + 2 * p - k) // s + 1
    dy = torch.randn(N, OC, oh, oh)
    dx_ref = ref.conv2d_backward_data(dy, w, (N, C, H, W), (s, s), (p, p))
    dx = ops.conv2d_backward_data(cl(dy), cl(w), (N, C, H, W), (s, s), (p, p))
    check_close(dx, dx_ref, _tol(OC), "conv dgrad")
    # accumulate into an existing buffer (how residual branch gradients are summed)
    base = torch.randn(N, C, H, W)
    out = cl(base.clone())
    ops.conv2d_backward_data(cl(dy), cl(w), (N, C, H, W), (s, s), (p, p), out=out, accumulate=True)
    check_close(out, dx_ref + base, _tol(OC), "conv dgrad accumulate")


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "x".join(map(str, c)))
def test_conv_backward_weight(case):
    N, C, H, W, OC, k, s, p, _ = case
    x = torch.randn(N, C, H, W)
    oh = (H + 2 * p - k) // s + 1
    dy = torch.randn(N, OC, oh, oh)
    dw_ref = torch.zeros(OC, C, k, k)
    ref.conv2d_backward_weight(dy, x, dw_ref, (s, s), (p, p))
    dw = cl(torch.zeros(OC, C, k, k))
    ops.conv2d_backward_weight(cl(dy), cl(x), dw, (s, s), (p, p))
    check_close(dw, dw_ref, _tol(C), "conv wgrad")
    ops.conv2d_backward_weight(cl(dy), cl(x), dw, (s, s), (p, p))  # beta=1: second call accumulates
    check_close(dw, 2 * dw_ref, _tol(C), "conv wgrad accumulate")
    db_ref = torch.zeros(OC, 1, 1, 1)
    ref.bias_grad(dy, db_ref)
    db = torch.zeros(OC, 1, 1, 1, device=DEV)
    ops.bias_grad(cl(dy), db)
    check_close(db, db_ref, 1e-4, "bias grad")


@pytest.mark.parametrize("case", [(8, 64, 16, 16, 128, 3, 1, 1), (4, 256, 8, 8, 64, 1, 1, 0), (6, 20, 9, 9, 12, 3, 2, 1)],
                         ids=lambda c: "x".join(map(str, c)))
def test_conv_wgrad_deterministic_mode(case, monkeypatch):
    """DCNN_DETERMINISTIC=1: split-K slabs summed in index order -- two runs give the same bits (and the right values),
    also when the call accumulates into an existing gradient; bias gradients likewise."""
    monkeypatch.setenv("DCNN_DETERMINISTIC", "1")
    N, C, H, W, OC, k, s, p = case
    x = torch.randn(N, C, H, W)
    oh = (H + 2 * p - k) // s + 1
    dy = torch.randn(N, OC, oh, oh)
    dw_ref = torch.zeros(OC, C, k, k)
    ref.conv2d_backward_weight(dy, x, dw_ref, (s, s), (p, p))
    xg, dyg = cl(x), cl(dy)
    runs = []
    for _ in range(3):
        dw = cl(torch.zeros(OC, C, k, k))
        ops.conv2d_backward_weight(dyg, xg, dw, (s, s), (p, p))
        ops.conv2d_backward_weight(dyg, xg, dw, (s, s), (p, p))
        db = torch.zeros(OC, 1, 1, 1, device=DEV)
        ops.bias_grad(dyg, db)
        runs.append((dw.clone(), db.clone()))
    check_close(runs[0][0], 2 * dw_ref, 4e-3, "deterministic wgrad")
    for dw, db in runs[1:]:
        assert torch.equal(dw, runs[0][0]) and torch.equal(db, runs[0][1])


@pytest.mark.parametrize("case", [(4, 64, 16, 16, 128, 3, 1, 1, True), (2, 16, 9, 9, 48, 5, 1, 0, True),
                                  (2, 3, 32, 32, 64, 3, 1, 1, False), (3, 6, 8, 8, 10, 3, 1, 1, True)],
                         ids=lambda c: "x".join(map(str, c)))
def test_conv_relu_epilogue_and_fused_bias_gradient(case):
    """Conv -> ReLU with the ReLU in the GEMM epilogue, and the bias gradient computed inside the weight-gradient launch
    (tensor-core paths, packed stem and the CUDA-core fallback) against the CPU oracle."""
    N, C, H, W, OC, k, s, p, bias = case
    x = torch.randn(N, C, H, W)
    w = torch.randn(OC, C, k, k) / (C * k * k) ** 0.5
    b = torch.randn(OC, 1, 1, 1) if bias else None
    y_ref, _ = ref.conv2d_forward(x, w, b, (s, s), (p, p), False)
    y, _ = ops.conv2d_forward(cl(x), cl(w), None if b is None else b.to(DEV), (s, s), (p, p), False, True)
    # after the ReLU a channel's largest value can be several times smaller than before it, so the per-channel bound
    # (relative to that maximum) needs more room for the same TF32 noise
    check_close(y, torch.relu(y_ref), 8e-3, "conv + relu epilogue")
    assert float(y.min()) >= 0.0
    dy = torch.randn_like(y_ref)
    dw_ref, db_ref = torch.zeros(OC, C, k, k), torch.zeros(OC, 1, 1, 1)
    ref.conv2d_backward_weight(dy, x, dw_ref, (s, s), (p, p), db_ref)
    dw, db = cl(torch.zeros(OC, C, k, k)), torch.zeros(OC, 1, 1, 1, device=DEV)
    ops.conv2d_backward_weight(cl(dy), cl(x), dw, (s, s), (p, p), db)
    ops.conv2d_backward_weight(cl(dy), cl(x), dw, (s, s), (p, p), db)       # both accumulate
    check_close(dw, 2 * dw_ref, 4e-3, "wgrad with fused bias gradient")
    check_close(db, 2 * db_ref, 4e-3, "fused bias gradient")


@pytest.mark.parametrize("B,Fin,Fout", [(256, 2048, 200), (33, 96, 48), (64, 512, 10)])
def test_dense_relu_epilogue_and_fused_bias_gradient(B, Fin, Fout):
    x, w, b = torch.randn(B, Fin), torch.randn(Fout, Fin) / Fin ** 0.5, torch.randn(Fout, 1, 1, 1)
    y = ops.dense_forward(x.to(DEV), w.to(DEV), b.to(DEV), True)
    check_close(y, torch.relu(ref.dense_forward(x, w, b)), 8e-3, "dense + relu epilogue")
    dy = torch.randn(B, Fout)
    dw_ref, db_ref = torch.zeros(Fout, Fin), torch.zeros(Fout, 1, 1, 1)
    ref.dense_backward_weight(dy, x, dw_ref, db_ref)
    dw, db = torch.zeros(Fout, Fin, device=DEV), torch.zeros(Fout, 1, 1, 1, device=DEV)
    ops.dense_backward_weight(dy.to(DEV), x.to(DEV), dw, db)
    check_close(dw, dw_ref, 4e-3, "dense wgrad")
    check_close(db, db_ref, 4e-3, "dense fused bias gradient")


def test_conv_fp32_exact_mode(monkeypatch):
    monkeypatch.setenv("DCNN_DISABLE_TC", "1")
    x, w = torch.randn(2, 64, 8, 8), torch.randn(32, 64, 3, 3) / 24
    y_ref, _ = ref.conv2d_forward(x, w, None, (1, 1), (1, 1))
    y, _ = ops.conv2d_forward(cl(x), cl(w), None, (1, 1), (1, 1))
    check_close(y, y_ref, 2e-5, "fp32 exact conv")


@pytest.mark.parametrize("B,Fin,Fout", [(256, 2048, 200), (64, 512, 10), (7, 96, 33), (32, 75, 10), (48, 100, 100),
                                        (300, 512, 1000)])
def test_dense(B, Fin, Fout):
    x, w, b = torch.randn(B, Fin), torch.randn(Fout, Fin) / Fin ** 0.5, torch.randn(Fout, 1, 1, 1)
    tol = 4e-3
    check_close(ops.dense_forward(x.to(DEV), w.to(DEV), b.to(DEV)), ref.dense_forward(x, w, b), tol, "dense fwd")
    dy = torch.randn(B, Fout)
    tol_d = 4e-3
    check_close(ops.dense_backward_data(dy.to(DEV), w.to(DEV)), ref.dense_backward_data(dy, w), tol_d, "dense dgrad")
    dw_ref = torch.zeros(Fout, Fin)
    ref.dense_backward_weight(dy, x, dw_ref)
    dw = torch.zeros(Fout, Fin, device=DEV)
    ops.dense_backward_weight(dy.to(DEV), x.to(DEV), dw)
    check_close(dw, dw_ref, tol, "dense wgrad")


def test_dense_forward_with_few_output_tiles_and_a_long_reduction():
    """One output tile and a 65536-long reduction: routed through the split-K weight-gradient kernel
    (ops/cuda.py::_dense_forward_split_k) instead of a forward launch that would occupy one SM."""
    M, N, K = 8, 12, 65536
    x, w, b = torch.randn(M, K), torch.randn(N, K) / K ** 0.5, torch.randn(N)
    y = ops.dense_forward(x.to(DEV), w.to(DEV), b.to(DEV), True)
    want = torch.relu(x @ w.t() + b)
    assert y.shape == want.shape
    assert float((y.cpu() - want).abs().max() / want.abs().max()) < 4e-3


# ----------------------------------------------------------------------------------------- batch norm
@pytest.mark.parametrize("N,C,H,W", [(8, 64, 16, 16), (4, 256, 8, 8), (3, 6, 5, 5), (16, 2048, 2, 2)])
@pytest.mark.parametrize("relu,res", [(False, False), (True, False), (True, True)])
def test_batchnorm_fwd_bwd(N, C, H, W, relu, res):
    x = torch.randn(N, C, H, W) * 3 + 5   # large mean: exercises the shifted/centred statistics
    gamma, beta = torch.rand(C, 1, 1, 1) + 0.5, torch.randn(C, 1, 1, 1)
    r = torch.randn(N, C, H, W) if res else None
    rm_ref, rv_ref = torch.zeros(C), torch.ones(C)
    mean_ref, inv_ref = ref.bn_finalize(ref.bn_stats(x), 1e-5, 0.1, rm_ref, rv_ref)
    y_ref = ref.bn_apply(x, mean_ref, inv_ref, gamma, beta, r, relu)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    xg, gg, bg = cl(x), gamma.to(DEV), beta.to(DEV)
    mean, inv = ops.bn_finalize(ops.bn_stats(xg), 1e-5, 0.1, rm, rv, gamma=gg, beta=bg)
    y = ops.bn_apply(xg, mean, inv, gg, bg, None if r is None else cl(r), relu)
    check_close(mean, mean_ref, 1e-5, "bn mean")
    check_close(inv[0], inv_ref, 1e-4, "bn invstd")
    check_close(rm, rm_ref, 1e-5, "running mean")
    check_close(rv, rv_ref, 1e-4, "running var")
    check_close(y, y_ref, 1e-4, "bn apply")
    dout = torch.randn(N, C, H, W)
    dg_ref, db_ref = torch.zeros(C, 1, 1, 1), torch.zeros(C, 1, 1, 1)
    dx_ref, m_ref = ref.bn_backward(dout, x, y_ref, mean_ref, inv_ref, gamma, dg_ref, db_ref, relu, True)
    dg, db = torch.zeros(C, 1, 1, 1, device=DEV), torch.zeros(C, 1, 1, 1, device=DEV)
    dx, m = ops.bn_backward(cl(dout), xg, y, mean, inv, gg, dg, db, relu, True)
    check_close(dx, dx_ref, 2e-4, "bn dx")
    check_close(m, m_ref, 1e-6, "masked dout")
    check_close(dg, dg_ref, 2e-4, "dgamma")
    check_close(db, db_ref, 2e-4, "dbeta")
    if relu and not res:
        # ReLU mask recomputed from x with the forward scale/shift: the saved output is not read at all
        dg2, db2 = torch.zeros(C, 1, 1, 1, device=DEV), torch.zeros(C, 1, 1, 1, device=DEV)
        dx2, m2 = ops.bn_backward(cl(dout), xg, None, mean, inv, gg, dg2, db2, relu, True, recompute_mask=True)
        check_close(m2, m, 1e-7, "masked dout (recomputed mask)")
        check_close(dx2, dx, 1e-6, "bn dx (recomputed mask)")
        check_close(dg2, dg, 1e-6, "dgamma (recomputed mask)")


def test_batchnorm_inference():
    x = torch.randn(4, 32, 6, 6)
    rm, rv, g, b = torch.randn(32), torch.rand(32) + 0.5, torch.rand(32, 1, 1, 1), torch.randn(32, 1, 1, 1)
    check_close(ops.bn_inference(cl(x), rm.to(DEV), rv.to(DEV), g.to(DEV), b.to(DEV), 1e-5, None, True),
                ref.bn_inference(x, rm, rv, g, b, 1e-5, None, True), 1e-5, "bn eval")


@pytest.mark.parametrize("N,C,H,W,G", [(4, 32, 8, 8, 8), (2, 6, 5, 5, 3), (3, 64, 4, 4, 32)])
def test_groupnorm(N, C, H, W, G):
    x = torch.randn(N, C, H, W) * 2 + 1
    gamma, beta = torch.rand(C, 1, 1, 1) + 0.5, torch.randn(C, 1, 1, 1)
    y_ref, mean_ref, inv_ref = ref.groupnorm_forward(x, G, gamma, beta, 1e-5)
    y, mean, inv = ops.groupnorm_forward(cl(x), G, gamma.to(DEV), beta.to(DEV), 1e-5)
    check_close(y, y_ref, 1e-4, "gn fwd")
    dy = torch.randn(N, C, H, W)
    dg_ref, db_ref = torch.zeros(C, 1, 1, 1), torch.zeros(C, 1, 1, 1)
    dx_ref = ref.groupnorm_backward(dy, x, mean_ref, inv_ref, G, gamma, dg_ref, db_ref)
    dg, db = torch.zeros(C, 1, 1, 1, device=DEV), torch.zeros(C, 1, 1, 1, device=DEV)
    dx = ops.groupnorm_backward(cl(dy), cl(x), mean, inv, G, gamma.to(DEV), dg, db)
    check_close(dx, dx_ref, 2e-4, "gn dx")
    check_close(dg, dg_ref, 2e-4, "gn dgamma")
    check_close(db, db_ref, 2e-4, "gn dbeta")


# ----------------------------------------------------------------------------------------- pooling / act / misc
@pytest.mark.parametrize("H", [13, 12])
@pytest.mark.parametrize("C", [64, 6])
@pytest.mark.parametrize("k,s,p", [(2, 2, 0), (3, 2, 1), (3, 3, 0), (3, 1, 1)])
def test_maxpool(C, k, s, p, H):
    x = torch.randn(3, C, H, H + 3)
    y_ref, idx_ref = ref.maxpool_forward(x, (k, k), (s, s), (p, p))
    y, idx = ops.maxpool_forward(cl(x), (k, k), (s, s), (p, p))
    check_close(y, y_ref, 1e-7, "maxpool fwd")
    dy = torch.randn_like(y_ref)
    dx_ref = ref._maxpool_bwd_scatter(dy, idx_ref, x.shape)
    dx = ops.maxpool_backward(cl(dy), idx, tuple(x.shape), (k, k), (s, s), (p, p))
    check_close(dx, dx_ref, 1e-6, "maxpool bwd")


@pytest.mark.parametrize("C", [64, 6])
@pytest.mark.parametrize("k,s,p", [(2, 2, 0), (4, 1, 0), (3, 2, 1), (4, 4, 0)])
def test_avgpool(C, k, s, p):
    x = torch.randn(3, C, 12, 12)
    y_ref = ref.avgpool_forward(x, (k, k), (s, s), (p, p))
    check_close(ops.avgpool_forward(cl(x), (k, k), (s, s), (p, p)), y_ref, 1e-6, "avgpool fwd")
    dy = torch.randn_like(y_ref)
    check_close(ops.avgpool_backward(cl(dy), tuple(x.shape), (k, k), (s, s), (p, p)),
                ref.avgpool_backward(dy, tuple(x.shape), (k, k), (s, s), (p, p)), 1e-6, "avgpool bwd")


@pytest.mark.parametrize("kind,alpha", [("relu", 0), ("leaky_relu", 0.01), ("elu", 1.0), ("sigmoid", 0), ("tanh", 0),
                                        ("softmax", 0), ("linear", 0)])
def test_activations(kind, alpha):
    x, g = torch.randn(5, 10, 3, 3), torch.randn(5, 10, 3, 3)
    check_close(ops.activation_forward(cl(x), kind, alpha), ref.activation_forward(x, kind, alpha), 1e-5, kind)
    check_close(ops.activation_backward(cl(x), cl(g), kind, alpha), ref.activation_backward(x, g, kind, alpha), 1e-5,
                kind + " grad")


def test_add_flatten_dropout():
    a, b = torch.randn(3, 5, 7, 7), torch.randn(3, 5, 7, 7)
    check_close(ops.add(cl(a), cl(b)), a + b, 1e-7, "add")
    f = ops.flatten_nchw(cl(a))
    assert torch.equal(f.cpu(), a.reshape(3, -1, 1, 1))
    assert torch.equal(ops.unflatten_nchw(f, a.shape).cpu(), a)
    y = ops.dropout_forward(torch.ones(1 << 16, device=DEV), 0.3, 7, 1)
    keep = float((y > 0).float().mean())
    assert abs(keep - 0.7) < 0.01 and abs(float(y.max()) - 1 / 0.7) < 1e-5
    assert torch.equal(y, ops.dropout_backward(torch.ones(1 << 16, device=DEV), 0.3, 7, 1))
    assert not torch.equal(y, ops.dropout_forward(torch.ones(1 << 16, device=DEV), 0.3, 7, 2))


@pytest.mark.parametrize("kind,param", [("softmax_crossentropy", 0), ("logsoftmax_crossentropy", 0),
                                        ("crossentropy", 1e-7), ("mse", 0), ("mae", 0), ("huber", 1.0)])
def test_losses(kind, param):
    B, C = 37, 200
    p = torch.randn(B, C, 1, 1)
    if kind == "crossentropy":
        p = torch.softmax(p, 1)
    t = F.one_hot(torch.randint(0, C, (B,)), C).float().reshape(B, C, 1, 1)
    l_ref, g_ref, c_ref = ref.loss_forward_backward(kind, p, t, param)
    l, g, c = ops.loss_forward_backward(kind, p.to(DEV), t.to(DEV), param)
    assert abs(float(l) - float(l_ref)) < 1e-4 * max(1.0, abs(float(l_ref)))
    check_close(g, g_ref, 1e-4, kind + " grad")
    if c_ref is not None:
        assert int(c) == int(c_ref)


@pytest.mark.parametrize("opt", ["sgd", "momentum", "adam", "adamw"])
def test_optimizers(opt):
    shapes = [(64, 3, 3, 3), (64, 1, 1, 1), (200, 2048, 1, 1), (7,), (100000,)]
    P_ = [torch.randn(s) for s in shapes]
    G_ = [torch.randn(s) for s in shapes]
    S1 = [torch.randn(s) * 0.1 for s in shapes]
    S2 = [torch.rand(s) * 0.1 for s in shapes]
    dev = [[t.clone().to(DEV) for t in L] for L in (P_, G_, S1, S2)]
    if opt in ("sgd", "momentum"):
        mom = 0.9 if opt == "momentum" else 0.0
        ref.sgd_update(P_, G_, S1, 0.1, mom)
        ops.sgd_update(dev[0], dev[1], dev[2], 0.1, mom, zero_grads=True)
    else:
        ref.adam_update(P_, G_, S1, S2, 1e-2, 0.9, 0.999, 1e-8, 0.1, 0.001, 5e-4, opt == "adamw")
        ops.adam_update(dev[0], dev[1], dev[2], dev[3], 1e-2, 0.9, 0.999, 1e-8, 0.1, 0.001, 5e-4, opt == "adamw",
                        zero_grads=True)
    for a, b in zip(dev[0], P_):
        check_close(a, b, 1e-5, opt + " params")
    for g in dev[1]:
        assert float(g.abs().max()) == 0.0


# ----------------------------------------------------------------------------------------- 256-pixel-tile conv kernel
def test_conv_tc3_large_problems():
    """conv_tc3 is only selected for loop spaces with >= 1 tile per SM; scripts/check_conv_v3.py runs such shapes
    (partial tiles, partial N tiles, strided boxes, accumulate mode) against cuDNN FP32 in a fresh process."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, DCNN_CONV_V3="1")
    r = subprocess.run([sys.executable, os.path.join(root, "scripts", "check_conv_v3.py")], env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ALL PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]
    assert "FAIL" not in r.stdout


# ----------------------------------------------------------------------------------------- kernel variants
@pytest.mark.parametrize("N,C,H,W", [(8, 64, 16, 16), (4, 256, 8, 8), (16, 2048, 2, 2)])
def test_batchnorm_relu_bitmask(N, C, H, W):
    """(BN + residual -> ReLU): the 4-bit-per-float4 mask written by the forward apply drives backward."""
    x = torch.randn(N, C, H, W) * 2 + 1
    gamma, beta = torch.rand(C, 1, 1, 1) + 0.5, torch.randn(C, 1, 1, 1)
    r = torch.randn(N, C, H, W)
    xg, gg, bg, rg = cl(x), gamma.to(DEV), beta.to(DEV), cl(r)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    mean, inv = ops.bn_finalize(ops.bn_stats(xg), 1e-5, 0.1, rm, rv, gamma=gg, beta=bg)
    y = ops.bn_apply(xg, mean, inv, gg, bg, rg, True)
    mask = torch.zeros(x.numel() // 4, dtype=torch.uint8, device=DEV)
    y2 = ops.bn_apply(xg, mean, inv, gg, bg, rg, True, mask_out=mask)
    assert torch.equal(y, y2)
    bits = (y.permute(0, 2, 3, 1).reshape(-1, 4) > 0).to(torch.uint8)
    want = bits[:, 0] | (bits[:, 1] << 1) | (bits[:, 2] << 2) | (bits[:, 3] << 3)
    assert torch.equal(mask, want)
    dout = cl(torch.randn(N, C, H, W))
    dg, db = torch.zeros(C, 1, 1, 1, device=DEV), torch.zeros(C, 1, 1, 1, device=DEV)
    dx, m = ops.bn_backward(dout, xg, y, mean, inv, gg, dg, db, True, True)
    dg2, db2 = torch.zeros_like(dg), torch.zeros_like(db)
    dx2, m2 = ops.bn_backward(dout, xg, None, mean, inv, gg, dg2, db2, True, True, relu_mask=mask)
    assert torch.equal(m, m2) and torch.equal(dx, dx2) and torch.equal(dg, dg2) and torch.equal(db, db2)


@pytest.mark.parametrize("case", [(4, 3, 32, 32, 64, 3, 1, 1, True), (2, 1, 28, 28, 8, 5, 1, 0, False),
                                  (3, 3, 17, 15, 32, 3, 2, 1, True)], ids=lambda c: "x".join(map(str, c)))
def test_stem_packed_taps(case):
    """Small-Cin convolutions as packed-tap im2col + 1x1 tensor-core conv (the default; DCNN_STEM_TC=0 switches it off)
    against the CPU oracle."""
    from dcnn_b200.ops import cuda as K
    N, C, H, W, OC, k, s, p, bias = case
    x = torch.randn(N, C, H, W)
    w = torch.randn(OC, C, k, k) / (C * k * k) ** 0.5
    b = torch.randn(OC, 1, 1, 1) if bias else None
    y_ref, _ = ref.conv2d_forward(x, w, b, (s, s), (p, p), False)
    oh, ow = y_ref.shape[2:]
    dy = torch.randn(N, OC, oh, ow)
    dw_ref = torch.zeros(OC, C, k, k)
    ref.conv2d_backward_weight(dy, x, dw_ref, (s, s), (p, p))
    old = K.STEM_TC
    K.STEM_TC = True
    try:
        assert K._stem_kp(C, k, k) == 32
        y, st = ops.conv2d_forward(cl(x), cl(w), None if b is None else b.to(DEV), (s, s), (p, p), True)
        check_close(y, y_ref, 4e-3, "stem fprop (packed taps)")
        mean, _ = ops.bn_finalize(st, 1e-5, 0.1, torch.zeros(OC, device=DEV), torch.ones(OC, device=DEV))
        assert (mean.cpu() - y_ref.mean((0, 2, 3))).abs().max() < 5e-3
        dw = cl(torch.zeros(OC, C, k, k))
        ops.conv2d_backward_weight(cl(dy), cl(x), dw, (s, s), (p, p))
        ops.conv2d_backward_weight(cl(dy), cl(x), dw, (s, s), (p, p))
        check_close(dw, 2 * dw_ref, 4e-3, "stem wgrad (packed taps, accumulates)")
    finally:
        K.STEM_TC = old


def test_bn_finalize_narrow(monkeypatch):
    """8-channel x 128-lane finalize kernels (the default from 256 statistics tiles on) == the wide ones
    (DCNN_BN_FINALIZE_NARROW=0)."""
    x = torch.randn(40, 64, 32, 32) * 2 + 3
    w = torch.randn(96, 64, 1, 1) / 8
    xg, wg = cl(x), cl(w)
    gamma, beta = (torch.rand(96, 1, 1, 1) + 0.5).to(DEV), torch.randn(96, 1, 1, 1).to(DEV)
    res = {}
    for narrow in ("0", "1"):
        monkeypatch.setenv("DCNN_BN_FINALIZE_NARROW", narrow)
        y, st = ops.conv2d_forward(xg, wg, None, (1, 1), (0, 0), True)
        assert st.tiles >= 256
        rm, rv = torch.zeros(96, device=DEV), torch.ones(96, device=DEV)
        mean, inv = ops.bn_finalize(st, 1e-5, 0.1, rm, rv, gamma=gamma, beta=beta)
        out = ops.bn_apply(y, mean, inv, gamma, beta, None, True)
        dg, db = torch.zeros(96, 1, 1, 1, device=DEV), torch.zeros(96, 1, 1, 1, device=DEV)
        dx, _ = ops.bn_backward(cl(torch.ones(40, 96, 32, 32)) * 0.5, y, out, mean, inv, gamma, dg, db, True, False)
        res[narrow] = [mean.clone(), inv[0].clone(), inv[1].clone(), inv[2].clone(), rm, rv, dg, db, dx]
    for a, b in zip(res["0"], res["1"]):
        check_close(b, a, 1e-6, "narrow finalize")


def test_conv_tc5_cta_pairs():
    """CTA-pair (tcgen05 cta_group::2) convolution kernel against cuDNN FP32 in a fresh process (DCNN_CONV_2CTA=1)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, DCNN_CONV_2CTA="1")
    r = subprocess.run([sys.executable, os.path.join(root, "scripts", "check_conv_v3.py"), "--pair"], env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ALL PASS" in r.stdout and "FAIL" not in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]


def test_conv_autotuner_picks_a_variant_and_stays_correct():
    """DCNN_AUTOTUNE: every structurally possible conv kernel variant is timed on first sight of a shape; results stay
    equal to the CPU oracle whichever wins, and the choice is cached."""
    from dcnn_b200.ops import cuda as K
    old = K.AUTOTUNE
    K.AUTOTUNE = True
    K._TUNED.clear()
    try:
        for (N, C, H, OC, k, p) in [(32, 64, 32, 128, 3, 1), (16, 128, 16, 256, 1, 0)]:
            x = torch.randn(N, C, H, H)
            w = torch.randn(OC, C, k, k) / (C * k * k) ** 0.5
            y_ref, st_ref = ref.conv2d_forward(x, w, None, (1, 1), (p, p), True)
            y, st = ops.conv2d_forward(cl(x), cl(w), None, (1, 1), (p, p), True)
            check_close(y, y_ref, 4e-3, "autotuned fprop")
            mean, _ = ops.bn_finalize(st, 1e-5, 0.1, torch.zeros(OC, device=DEV), torch.ones(OC, device=DEV))
            assert (mean.cpu() - y_ref.mean((0, 2, 3))).abs().max() < 5e-3
            key = ("f", N, H, H, C, OC, k, k, 1, 1)
            assert K._TUNED.get(key) in (2, 3, 5)
            y2, _ = ops.conv2d_forward(cl(x), cl(w), None, (1, 1), (p, p), True)      # cached choice, same answer
            check_close(y2, y_ref, 4e-3, "autotuned fprop (cached)")
            dy = torch.randn_like(y_ref)
            dx_ref = ref.conv2d_backward_data(dy, w, (N, C, H, H), (1, 1), (p, p))
            dx = ops.conv2d_backward_data(cl(dy), cl(w), (N, C, H, H), (1, 1), (p, p))
            check_close(dx, dx_ref, 4e-3, "autotuned dgrad")
    finally:
        K.AUTOTUNE = old
        K._TUNED.clear()
