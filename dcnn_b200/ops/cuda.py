"""CUDA backend of the functional op set: thin wrappers that allocate outputs with torch and launch the
hand-written sm_100a kernels of libdcnn_b200.so on the current torch stream.  No cuDNN / cuBLAS / ATen compute
kernel runs on these paths (torch is used for allocation, views and stream/graph plumbing only).

Layout contract: 4-D activations are logical NCHW stored channels-last (NHWC); conv weights are logical OIHW
stored O-H-W-I.  Tensor-core (tcgen05) kernels are used whenever the GEMM-K channel count is a multiple of 32;
everything else takes the exact-FP32 CUDA-core kernels.  ``DCNN_DISABLE_TC=1`` forces the CUDA-core path (the
"FP32-exact mode" for bit-tight tests)."""
from __future__ import annotations

import os
from typing import Optional

import torch

from .._lib import D, F, I, LL, P, ULL, check, lib

NAME = "cuda"
CL = torch.channels_last

_ACT = {"relu": 0, "leaky_relu": 1, "elu": 2, "sigmoid": 3, "tanh": 4, "linear": 5}


def _stream():
    return P(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]):
    return P(t.data_ptr()) if t is not None else P(0)


# DCNN_BN_RECOMPUTE_MASK=0 restores the read-the-saved-output ReLU mask in BatchNorm backward
RECOMPUTE_RELU_MASK = os.environ.get("DCNN_BN_RECOMPUTE_MASK", "1") != "0"
# DCNN_BN_BITMASK=1: (BN + residual -> ReLU) units keep a 4-bit-per-float4 mask instead of re-reading their output in
# backward (opt-in until its hardware validation run)
RELU_BITMASK = os.environ.get("DCNN_BN_BITMASK", "1") == "1"


# DCNN_DETERMINISTIC=1: bit-reproducible gradients -- the split-K partial sums of the tensor-core weight gradient go to
# per-split slabs that a second kernel adds in index order (instead of red.global.add), and the bias-gradient column sums
# use one CTA per column group.  Costs one extra pass over splits x |dW| floats per convolution.
def deterministic() -> bool:
    return os.environ.get("DCNN_DETERMINISTIC", "0") == "1"


def _wgrad_scratch(N, OH, OW, K, Cin, sh_, sw_, ntaps, ktot, device):
    if not deterministic():
        return None
    splits = lib().dcnn_wgrad_tc_splits(I(N), I(OH), I(OW), I(K), I(Cin), I(sh_), I(sw_), I(ntaps))
    return torch.empty(splits * K * ktot, device=device, dtype=torch.float32)


def tc_enabled() -> bool:
    return os.environ.get("DCNN_DISABLE_TC", "0") != "1"


def tc_strided_enabled() -> bool:
    return os.environ.get("DCNN_DISABLE_TC_STRIDED", "0") != "1"


def as_nhwc(x: torch.Tensor) -> torch.Tensor:
    if x.dtype != torch.float32:
        x = x.float()
    return x.contiguous(memory_format=CL)


def _empty_nhwc(n, c, h, w, device):
    return torch.empty((n, c, h, w), device=device, dtype=torch.float32, memory_format=CL)


_TAPS = {}


def _taps(key, gen):
    t = _TAPS.get(key)
    if t is None:
        dh, dw, wi = gen()
        n = len(dh)
        t = (n, (I * n)(*dh), (I * n)(*dw), (I * n)(*wi))
        _TAPS[key] = t
    return t


class CudaStats:
    """BN partial statistics: ``partials`` = [tiles][2][C] (sum, centred M2) followed by [tiles] row counts."""
    __slots__ = ("partials", "tiles", "C", "count")

    def __init__(self, partials, tiles, C_, count):
        self.partials, self.tiles, self.C, self.count = partials, tiles, C_, count


# ================================================================================== convolution
# Small-Cin convolutions run as packed-tap im2col + a 1x1 tensor-core convolution instead of on CUDA cores: always for
# channel counts that are not a multiple of 4 (RGB / grey-scale stems: the 3x3x3 ResNet stem is K = 27 -> 32, the 7x7x3
# ImageNet stem K = 147 -> 160), and for other counts below 32 when all taps fit in 64 packed channels (cheaper than
# padding every tap to 32).  DCNN_STEM_TC=0 sends them to the CUDA-core kernel.
STEM_TC = os.environ.get("DCNN_STEM_TC", "1") == "1"


def _stem_kp(Cin, R, S):
    """Padded reduction length if this convolution qualifies for the packed-tap path, else 0."""
    if not (STEM_TC and tc_enabled()) or Cin % 32 == 0:
        return 0
    K = R * S * Cin
    limit = 64 if Cin % 4 == 0 else 256
    return (K + 31) // 32 * 32 if K <= limit else 0


# The packed activations of the most recent packed-tap forward call, for the weight gradient of the same layer:
# (input tensor object, its version, geometry, packed tensor).  Keyed on the tensor OBJECT (kept alive here), not its
# address -- a fresh batch may land at the address of the previous one with the same version number.  One entry: a
# training step runs the stem's backward after its own forward; anything else (microbatched pipelines) misses and packs
# again.  Assigned and read as one tuple, so stage threads can only miss, never see a mixed entry.
_PACK_CACHE = None


def _pack_taps_cached(x_orig, R, S, stride, pad, OH, OW, Kp, store):
    global _PACK_CACHE
    key = (R, S, tuple(stride), tuple(pad), OH, OW, Kp)
    ent = _PACK_CACHE
    if not store and ent is not None and ent[0] is x_orig and ent[1] == x_orig._version and ent[2] == key:
        _PACK_CACHE = None       # used once: the 134 MB of a ResNet stem at batch 256 go back to the allocator
        return ent[3]
    packed = _pack_taps(as_nhwc(x_orig), R, S, stride, pad, OH, OW, Kp)
    _PACK_CACHE = (x_orig, x_orig._version, key, packed) if store else None
    return packed


def _pack_taps(x, R, S, stride, pad, OH, OW, Kp):
    N, Cin, H, W = x.shape
    P_ = _empty_nhwc(N, Kp, OH, OW, x.device)
    check(lib().dcnn_pack_taps(_p(x), _p(P_), I(N), I(H), I(W), I(Cin), I(OH), I(OW), I(R), I(S), I(stride[0]),
                               I(stride[1]), I(pad[0]), I(pad[1]), I(Kp), _stream()), "pack_taps")
    return P_


def _packed_weights(w, Kp):
    """[Cout][R][S][Cin] (the stored O-H-W-I layout) -> [Cout][Kp] zero padded, as a 1x1 O-H-W-I weight."""
    Cout, Cin, R, S = w.shape
    flat = w.permute(0, 2, 3, 1).reshape(Cout, R * S * Cin)
    wp = torch.zeros(Cout, Kp, device=w.device, dtype=torch.float32)
    wp[:, :R * S * Cin] = flat
    return wp.view(Cout, Kp, 1, 1).contiguous(memory_format=CL)


# DCNN_AUTOTUNE=1: the first time a convolution shape is seen (outside CUDA-graph capture) every structurally possible
# kernel variant -- conv_tc2 (128-pixel tiles), conv_tc3 (256-pixel tiles), conv_tc5 (CTA pairs, Cout >= 256) -- is timed
# with CUDA events on that very call's operands and the fastest is used from then on, instead of the built-in size
# rules.  Validated on hardware (-0.18 ms on the 16.5 ms ResNet-50 step, profiles/05); opt-in because the timing run
# synchronises the host, which a pipeline stage with pre-posted receives must not do outside its warm-up.
# DCNN_AUTOTUNE_LOG=<file> appends one line per decision.
AUTOTUNE = os.environ.get("DCNN_AUTOTUNE", "0") == "1"
_TUNED = {}


def _autotuned(key, cout, fn, tunable=True):
    """Run ``fn()`` (one convolution launch sequence, idempotent when ``tunable``) under the variant chosen for ``key``."""
    if not AUTOTUNE:
        return fn()
    v = _TUNED.get(key)
    if v is None:
        if not tunable or torch.cuda.is_current_stream_capturing():
            return fn()
        best_t, times = None, []
        for cand in (2, 3, 5) if cout >= 256 else (2, 3):
            check(lib().dcnn_conv_tc_set_variant(I(cand)), "conv_tc_set_variant")
            try:
                fn()                                                 # warm (attribute set-up, descriptor caches)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3):
                    fn()
                e1.record()
                e1.synchronize()
                t = e0.elapsed_time(e1)
            finally:
                check(lib().dcnn_conv_tc_set_variant(I(0)), "conv_tc_set_variant")
            times.append((cand, round(t / 3 * 1000.0, 1)))
            if best_t is None or t < best_t:
                best_t, v = t, cand
        _TUNED[key] = v
        log = os.environ.get("DCNN_AUTOTUNE_LOG", "")
        if log:
            with open(log, "a") as f:
                f.write(f"{key} -> conv_tc{v}  us per launch {times}\n")
    check(lib().dcnn_conv_tc_set_variant(I(v)), "conv_tc_set_variant")
    try:
        return fn()
    finally:
        check(lib().dcnn_conv_tc_set_variant(I(0)), "conv_tc_set_variant")


def _conv_tc(x, N, H, W, Cin, w, Cout, Ktot, out, out_off, sn, sh, sw, OH, OW, mul_h, mul_w, taps, bias,
             accumulate, partials, relu=False):
    n, dh, dw, wi = taps
    check(lib().dcnn_conv_tc(_p(x), I(N), I(H), I(W), I(Cin), _p(w), I(Cout), I(Ktot),
                             P(out.data_ptr() + 4 * out_off), LL(sn), LL(sh), LL(sw), I(OH), I(OW), I(mul_h),
                             I(mul_w), I(n), dh, dw, wi, _p(bias), I(1 if accumulate else 0), I(1 if relu else 0),
                             _p(partials), _stream()), "dcnn_conv_tc")


def conv2d_forward(x, w, bias, stride, pad, want_stats=False, relu=False):
    """``relu``: max(., 0) in the epilogue of the tensor-core kernels (Conv -> ReLU without BatchNorm)."""
    assert not (relu and want_stats)
    x_orig = x
    N, Cin, H, W = x.shape
    Cout, _, R, S = w.shape
    assert w.is_contiguous(memory_format=CL), "conv weights must be stored O-H-W-I (channels_last)"
    sh_, sw_ = stride
    ph, pw = pad
    OH = (H + 2 * ph - R) // sh_ + 1
    OW = (W + 2 * pw - S) // sw_ + 1
    kp = _stem_kp(Cin, R, S)
    if kp:
        packed = _pack_taps_cached(x_orig, R, S, stride, pad, OH, OW, kp, store=True)
        return conv2d_forward(packed, _packed_weights(w, kp), bias, (1, 1), (0, 0), want_stats, relu)
    x = as_nhwc(x)
    y = _empty_nhwc(N, Cout, OH, OW, x.device)
    # tensor cores for every channel count that keeps rows 16-byte aligned: the TMA unit pads the reduction dimension
    # to a multiple of 32 with zeros (C = 1 / 3 stems take the packed-tap route above; anything else is CUDA cores)
    use_tc = tc_enabled() and Cin % 4 == 0 and Cout % 4 == 0 and R * S <= 49 and \
        ((sh_ == 1 and sw_ == 1) or (tc_strided_enabled() and sh_ <= 8 and sw_ <= 8))
    if use_tc:
        taps = _taps(("f", R, S, ph, pw), lambda: (
            [r - ph for r in range(R) for s in range(S)],
            [s - pw for r in range(R) for s in range(S)],
            [r * S + s for r in range(R) for s in range(S)]))
        def run():
            stats = None
            partials = None
            if want_stats:      # the statistics-tile count depends on the kernel variant in force
                tiles = lib().dcnn_conv_tc_stat_tiles(I(N), I(OH), I(OW), I(sh_), I(sw_), I(Cout),
                                                      I(R * S * ((Cin + 31) // 32)))
                partials = torch.empty(tiles * (2 * Cout + 1), device=x.device, dtype=torch.float32)
                stats = CudaStats(partials, tiles, Cout, N * OH * OW)
            _conv_tc(x, N, H, W, Cin, w, Cout, R * S * Cin, y, 0, OH * OW * Cout, OW * Cout, Cout, OH, OW, sh_, sw_,
                     taps, bias, False, partials, relu)
            return y, stats
        return _autotuned(("f", N, H, W, Cin, Cout, R, S, sh_, sw_), Cout, run)
    check(lib().dcnn_conv_fprop_simt(_p(x), _p(w), _p(bias), _p(y), I(N), I(H), I(W), I(Cin), I(Cout), I(R), I(S),
                                     I(sh_), I(sw_), I(ph), I(pw), _stream()), "dcnn_conv_fprop_simt")
    if relu:
        y = activation_forward(y, "relu", 0.0)
    return y, (bn_stats(y) if want_stats else None)


# The dgrad operand layout (W[K][R][S][C] -> Wt[C][R][S][K], spatial axes flipped) of EVERY weight tensor that has been
# through a backward pass is rebuilt by ONE batched launch, the first time a backward pass needs any of them after the
# weights may have changed.  An entry is valid while (a) torch has not modified the tensor in place (`_version`) and
# (b) none of this library's kernels that write parameters through raw pointers -- optimizer updates, the fused
# data-parallel update, CUDA-graph replays of either -- has run since (`weights_changed()` bumps the generation).
# Before: one launch per convolution per backward pass (52 per ResNet-50 step, 0.29 ms); microbatched and pipelined
# steps, which run many backward passes per update, reuse the layouts outright.
class _WtEntry:
    __slots__ = ("w", "wt", "stamp", "shape", "used")


_WT_CACHE = {}
_WT_GENERATION = [0]

def weights_changed():
    _WT_GENERATION[0] += 1


def clear_weight_layout_cache():
    _WT_CACHE.clear()


def _wt_stamp(w):
    return (w._version, _WT_GENERATION[0], torch.cuda.is_current_stream_capturing())


def _dgrad_weights(w):
    """W[K][R][S][C] -> Wt[C][R][S][K], both spatial axes flipped (K-major over the dY channels)."""
    K, Cc, R, S = w.shape
    key = (w.data_ptr(), K, Cc, R, S, w.device.index)
    stamp = _wt_stamp(w)
    ent = _WT_CACHE.get(key)
    if ent is not None:
        ent.used = _WT_GENERATION[0]
        if ent.stamp == stamp:
            return ent.wt
    if ent is None:
        if len(_WT_CACHE) > 4096:
            clear_weight_layout_cache()
        ent = _WtEntry()
        ent.w, ent.shape, ent.wt, ent.stamp, ent.used = w, (K, R, S, Cc), None, None, _WT_GENERATION[0]
        _WT_CACHE[key] = ent
    capturing = stamp[2]
    # everything on this device that is stale and was needed recently (a model that is no longer trained drops out)
    # goes into the same launch
    recent = _WT_GENERATION[0] - 4
    todo = [e for e in _WT_CACHE.values()
            if e is ent or (e.w.device == w.device and e.used >= recent and e.stamp != _wt_stamp(e.w))]
    rows = []
    for e in todo:
        # a buffer that belongs to a captured graph's pool is never written by eager launches (and vice versa)
        if e.wt is None or e.stamp is None or e.stamp[2] != capturing:
            k_, r_, s_, c_ = e.shape
            e.wt = torch.empty(c_ * r_ * s_ * k_, device=w.device, dtype=torch.float32)
        k_, r_, s_, c_ = e.shape
        rows.append([e.w.data_ptr(), e.wt.data_ptr(), k_, r_, s_, c_])
    if len(rows) == 1:
        k_, r_, s_, c_ = ent.shape
        check(lib().dcnn_weight_dgrad_layout(_p(ent.w), _p(ent.wt), I(k_), I(r_), I(s_), I(c_), _stream()),
              "weight_dgrad_layout")
    else:
        flat = [v for row in rows for v in row]
        check(lib().dcnn_weight_dgrad_layout_batched((LL * len(flat))(*flat), I(len(rows)), _stream()),
              "weight_dgrad_layout_batched")
    for e in todo:
        e.stamp = _wt_stamp(e.w)
    return ent.wt


def conv2d_backward_data(dy, w, x_shape, stride, pad, out=None, accumulate=False):
    dy = as_nhwc(dy)
    N, Cin, H, W = x_shape
    K, _, R, S = w.shape
    _, _, OH, OW = dy.shape
    sh_, sw_ = stride
    ph, pw = pad
    if out is None:
        accumulate = False
    dx = out if out is not None else _empty_nhwc(N, Cin, H, W, dy.device)
    assert dx.is_contiguous(memory_format=CL)
    use_tc = tc_enabled() and K % 4 == 0 and Cin % 4 == 0 and R * S <= 49 and R - 1 - ph >= -64
    if use_tc and sh_ == 1 and sw_ == 1:
        wt = _dgrad_weights(w)
        taps = _taps(("d1", R, S, ph, pw), lambda: (
            [r2 - (R - 1 - ph) for r2 in range(R) for s2 in range(S)],
            [s2 - (S - 1 - pw) for r2 in range(R) for s2 in range(S)],
            [r2 * S + s2 for r2 in range(R) for s2 in range(S)]))
        def run():
            _conv_tc(dy, N, OH, OW, K, wt, Cin, R * S * K, dx, 0, H * W * Cin, W * Cin, Cin, H, W, 1, 1, taps, None,
                     accumulate, None)
            return dx
        # an accumulating call is not idempotent: it uses the variant tuned by a plain call of the same shape, if any
        return _autotuned(("d", N, H, W, Cin, K, R, S), Cin, run, tunable=not accumulate)
    if use_tc:
        # strided dgrad = one stride-1 tap-subset convolution per output parity class, written with an output stride
        wt = _dgrad_weights(w)
        classes = []
        empty_class = False
        for a in range(sh_):
            for b in range(sw_):
                if a >= H or b >= W:
                    continue
                rs = [r for r in range(R) if (a + ph - r) % sh_ == 0]
                ss = [s for s in range(S) if (b + pw - s) % sw_ == 0]
                if not rs or not ss:
                    empty_class = True
                    continue
                classes.append((a, b, rs, ss))
        if empty_class and not accumulate:
            dx.zero_()
        for a, b, rs, ss in classes:
            taps = _taps(("ds", R, S, ph, pw, sh_, sw_, a, b), lambda: (
                [(a + ph - r) // sh_ for r in rs for s in ss],
                [(b + pw - s) // sw_ for r in rs for s in ss],
                [(R - 1 - r) * S + (S - 1 - s) for r in rs for s in ss]))
            Hc = (H - a + sh_ - 1) // sh_
            Wc = (W - b + sw_ - 1) // sw_
            _conv_tc(dy, N, OH, OW, K, wt, Cin, R * S * K, dx, (a * W + b) * Cin, H * W * Cin, sh_ * W * Cin,
                     sw_ * Cin, Hc, Wc, 1, 1, taps, None, accumulate, None)
        return dx
    check(lib().dcnn_conv_dgrad_simt(_p(dy), _p(w), _p(dx), I(N), I(H), I(W), I(Cin), I(K), I(R), I(S), I(sh_),
                                     I(sw_), I(ph), I(pw), I(1 if accumulate else 0), _stream()), "conv_dgrad_simt")
    return dx


def conv2d_backward_weight(dy, x, dw, stride, pad, db=None):
    """dw += dL/dW; with ``db`` also db += column sums of dy -- inside the same launch on the tensor-core path (one extra
    N = 16 MMA against a tile of ones), by the column-sum kernel otherwise."""
    dy = as_nhwc(dy)
    x_orig = x
    N, Cin, H, W = x.shape
    K, _, R, S = dw.shape
    _, _, OH, OW = dy.shape
    sh_, sw_ = stride
    ph, pw = pad
    assert dw.is_contiguous(memory_format=CL)
    kp = _stem_kp(Cin, R, S)
    if kp and K % 4 == 0:
        # the packed activations of this layer's forward call are reused when they are still the latest (one 53 us
        # pack_taps launch and one layout conversion less per ResNet-50 step); otherwise they are rebuilt
        dwp = torch.zeros(K, kp, 1, 1, device=dw.device, dtype=torch.float32).contiguous(memory_format=CL)
        packed = _pack_taps_cached(x_orig, R, S, stride, pad, OH, OW, kp, store=False)
        conv2d_backward_weight(dy, packed, dwp, (1, 1), (0, 0), db)
        dw.permute(0, 2, 3, 1).reshape(K, R * S * Cin).add_(dwp.reshape(K, kp)[:, :R * S * Cin])
        return
    x = as_nhwc(x)
    use_tc = tc_enabled() and Cin % 4 == 0 and K % 4 == 0 and R * S <= 49 and \
        ((sh_ == 1 and sw_ == 1) or (tc_strided_enabled() and sh_ <= 8 and sw_ <= 8))
    if use_tc:
        taps = _taps(("f", R, S, ph, pw), lambda: (
            [r - ph for r in range(R) for s in range(S)],
            [s - pw for r in range(R) for s in range(S)],
            [r * S + s for r in range(R) for s in range(S)]))
        n, dh, dw_, wi = taps
        scratch = _wgrad_scratch(N, OH, OW, K, Cin, sh_, sw_, n, R * S * Cin, dy.device)
        fused_db = db if scratch is None else None
        check(lib().dcnn_wgrad_tc(_p(dy), I(N), I(OH), I(OW), I(K), _p(x), I(H), I(W), I(Cin), _p(dw),
                                  I(R * S * Cin), I(sh_), I(sw_), I(n), dh, dw_, wi, _p(scratch), _p(fused_db),
                                  _stream()), "dcnn_wgrad_tc")
        if db is not None and fused_db is None:
            bias_grad(dy, db)
        return
    if db is not None:
        bias_grad(dy, db)
    check(lib().dcnn_conv_wgrad_simt(_p(dy), _p(x), _p(dw), I(N), I(H), I(W), I(Cin), I(K), I(R), I(S), I(sh_),
                                     I(sw_), I(ph), I(pw), _stream()), "conv_wgrad_simt")


def bias_grad(dy, db):
    dy = as_nhwc(dy) if dy.dim() == 4 else dy.contiguous()
    Cc = dy.shape[1]
    M = dy.numel() // Cc
    check(lib().dcnn_colsum_accum(_p(dy), LL(M), I(Cc), _p(db), I(1 if deterministic() else 0), _stream()),
          "colsum_accum")


# ================================================================================== dense (1x1 conv on [B,1,1,C])
_ONE_TAP = None


def _one_tap():
    global _ONE_TAP
    if _ONE_TAP is None:
        _ONE_TAP = (1, (I * 1)(0), (I * 1)(0), (I * 1)(0))
    return _ONE_TAP


def _reduction_bound(B, Fout, Fin):
    """A product with a handful of output tiles and a very long reduction (the reference's gemm_benchmark shape,
    64 x 128 x 2M): the forward kernel has no split-K, so it would run on as many SMs as there are output tiles."""
    return (B + 127) // 128 * ((Fout + 127) // 128) <= 8 and Fin >= 32768 and B % 4 == 0


def _dense_forward_split_k(x2d, w2d, bias, relu):
    """y = x w^T through the weight-gradient kernel, whose reduction dimension is split over the whole machine:
    dW'[B][Fout] = (x^T)^T (w^T) with the Fin reduction rows as its 'pixels' (two transposed copies, then one launch)."""
    B, Fout = x2d.shape[0], w2d.shape[0]
    y = torch.zeros((B, Fout), device=x2d.device, dtype=torch.float32)
    dense_backward_weight(x2d.t().contiguous(), w2d.t().contiguous(), y)
    if bias is not None:
        y += bias.reshape(1, Fout)
    return torch.relu_(y) if relu else y


def dense_forward(x2d, w2d, bias, relu=False):
    x2d = x2d.contiguous()
    B, Fin = x2d.shape
    Fout = w2d.shape[0]
    if tc_enabled() and Fin % 4 == 0 and Fout % 4 == 0 and _reduction_bound(B, Fout, Fin):
        return _dense_forward_split_k(x2d, w2d.reshape(Fout, Fin), bias, relu)
    y = torch.empty((B, Fout), device=x2d.device, dtype=torch.float32)
    if tc_enabled() and Fin % 4 == 0 and Fout % 4 == 0:
        _conv_tc(x2d, B, 1, 1, Fin, w2d, Fout, Fin, y, 0, Fout, Fout, Fout, 1, 1, 1, 1, _one_tap(), bias, False, None,
                 relu)
        return y
    else:
        check(lib().dcnn_conv_fprop_simt(_p(x2d), _p(w2d), _p(bias), _p(y), I(B), I(1), I(1), I(Fin), I(Fout), I(1),
                                         I(1), I(1), I(1), I(0), I(0), _stream()), "dense fprop simt")
    if relu:
        y = activation_forward(y.view(B, Fout, 1, 1), "relu", 0.0).view(B, Fout)
    return y


def dense_backward_data(dy2d, w2d):
    dy2d = dy2d.contiguous()
    B, Fout = dy2d.shape
    Fin = w2d.shape[1]
    dx = torch.empty((B, Fin), device=dy2d.device, dtype=torch.float32)
    if tc_enabled() and Fout % 4 == 0 and Fin % 4 == 0:
        wt = _dgrad_weights(w2d.view(Fout, Fin, 1, 1))
        _conv_tc(dy2d, B, 1, 1, Fout, wt, Fin, Fout, dx, 0, Fin, Fin, Fin, 1, 1, 1, 1, _one_tap(), None, False, None)
    else:
        check(lib().dcnn_conv_dgrad_simt(_p(dy2d), _p(w2d), _p(dx), I(B), I(1), I(1), I(Fin), I(Fout), I(1), I(1),
                                         I(1), I(1), I(0), I(0), I(0), _stream()), "dense dgrad simt")
    return dx


def dense_backward_weight(dy2d, x2d, dw2d, db=None):
    dy2d = dy2d.contiguous()
    x2d = x2d.contiguous()
    B, Fout = dy2d.shape
    Fin = x2d.shape[1]
    if tc_enabled() and Fin % 4 == 0 and Fout % 4 == 0:
        n, dh, dw_, wi = _one_tap()
        scratch = _wgrad_scratch(B, 1, 1, Fout, Fin, 1, 1, 1, Fin, dy2d.device)
        fused_db = db if scratch is None else None
        check(lib().dcnn_wgrad_tc(_p(dy2d), I(B), I(1), I(1), I(Fout), _p(x2d), I(1), I(1), I(Fin), _p(dw2d), I(Fin),
                                  I(1), I(1), I(n), dh, dw_, wi, _p(scratch), _p(fused_db), _stream()), "dense wgrad tc")
        if db is not None and fused_db is None:
            bias_grad(dy2d.view(B, Fout, 1, 1), db)
    else:
        if db is not None:
            bias_grad(dy2d.view(B, Fout, 1, 1), db)
        check(lib().dcnn_conv_wgrad_simt(_p(dy2d), _p(x2d), _p(dw2d), I(B), I(1), I(1), I(Fin), I(Fout), I(1), I(1),
                                         I(1), I(1), I(0), I(0), _stream()), "dense wgrad simt")


# ================================================================================== batch norm
def bn_stats(x):
    x = as_nhwc(x)
    Cc = x.shape[1]
    M = x.numel() // Cc
    tiles = lib().dcnn_bn_stat_tiles(LL(M))
    partials = torch.empty(tiles * (2 * Cc + 1), device=x.device, dtype=torch.float32)
    check(lib().dcnn_bn_colstats(_p(x), LL(M), I(Cc), _p(partials), _stream()), "bn_colstats")
    return CudaStats(partials, tiles, Cc, M)


class _BNCoeffs:
    """mean / invstd plus the fused scale/shift used by the apply pass."""
    __slots__ = ("mean", "invstd", "scale", "shift")


def bn_finalize(stats, eps, momentum, running_mean, running_var, sync=None, gamma=None, beta=None):
    Cc = stats.C
    dev = stats.partials.device
    buf = torch.empty(4 * Cc, device=dev, dtype=torch.float32)
    mean, invstd, scale, shift = buf[:Cc], buf[Cc:2 * Cc], buf[2 * Cc:3 * Cc], buf[3 * Cc:]
    if sync is None:
        check(lib().dcnn_bn_finalize(_p(stats.partials), I(stats.tiles), I(Cc), F(eps), F(momentum),
                                     _p(running_mean), _p(running_var), _p(gamma), _p(beta), _p(mean), _p(invstd),
                                     _p(scale), _p(shift), _stream()), "bn_finalize")
    else:
        sync.finalize_forward(stats, eps, momentum, running_mean, running_var, gamma, beta, mean, invstd, scale, shift)
    return mean, (invstd, scale, shift)


def bn_apply(x, mean, invstd, gamma, beta, residual=None, relu=False, mask_out=None):
    """``mask_out`` (uint8, numel/4, ReLU only): also emit the compact ReLU mask for :func:`bn_backward`."""
    x = as_nhwc(x)
    if isinstance(invstd, tuple):
        invstd, scale, shift = invstd
    else:  # raw tensors: build the coefficients here
        scale = invstd if gamma is None else invstd * gamma.view(-1)
        shift = -mean * scale + (beta.view(-1) if beta is not None else 0)
    Cc = x.shape[1]
    M = x.numel() // Cc
    y = torch.empty_like(x, memory_format=torch.preserve_format)
    if residual is not None:
        residual = as_nhwc(residual)
    if mask_out is not None:
        assert relu and Cc % 4 == 0 and mask_out.dtype == torch.uint8 and mask_out.numel() * 4 == x.numel()
        check(lib().dcnn_bn_apply_mask(_p(x), _p(scale), _p(shift), _p(residual), _p(y), _p(mask_out), LL(M), I(Cc),
                                       _stream()), "bn_apply_mask")
        return y
    check(lib().dcnn_bn_apply(_p(x), _p(scale), _p(shift), _p(residual), _p(y), LL(M), I(Cc), I(1 if relu else 0),
                              _stream()), "bn_apply")
    return y


def bn_backward(dout, x, out, mean, invstd, gamma, dgamma, dbeta, relu=False, want_masked=False, sync=None,
                recompute_mask=False, relu_mask=None):
    """``recompute_mask`` (BN -> ReLU with no residual in between): the ReLU mask is re-evaluated from ``x`` with the
    forward pass's scale/shift instead of reading the saved output -- two activation-sized reads fewer."""
    dout = as_nhwc(dout)
    relu_mode = 1 if relu else 0
    if isinstance(invstd, tuple):
        if relu and recompute_mask and RECOMPUTE_RELU_MASK:
            scale, shift = invstd[1], invstd[2]
            if scale.data_ptr() + 4 * scale.numel() == shift.data_ptr():
                relu_mode, out = 2, scale       # kernels read scale[C] then shift[C] through this pointer
        invstd = invstd[0]
    if relu and relu_mask is not None and relu_mode == 1:
        relu_mode, out = 3, relu_mask          # 4 bits per float4 written by bn_apply(mask_out=...)
    Cc = x.shape[1]
    M = x.numel() // Cc
    tiles = lib().dcnn_bn_bwd_tiles(LL(M), I(Cc))
    partials = torch.empty(tiles * 2 * Cc, device=x.device, dtype=torch.float32)
    check(lib().dcnn_bn_bwd_reduce(_p(dout), _p(x), _p(out), _p(mean), _p(invstd), LL(M), I(Cc),
                                   I(relu_mode), _p(partials), _stream()), "bn_bwd_reduce")
    k = torch.empty(3 * Cc, device=x.device, dtype=torch.float32)
    k1, k2, k3 = k[:Cc], k[Cc:2 * Cc], k[2 * Cc:]
    if sync is None:
        check(lib().dcnn_bn_bwd_finalize(_p(partials), I(tiles), I(Cc), D(float(M)), _p(mean), _p(invstd), _p(gamma),
                                         _p(dgamma), _p(dbeta), _p(k1), _p(k2), _p(k3), _stream()),
              "bn_bwd_finalize")
    else:
        sync.finalize_backward(partials, tiles, Cc, M, mean, invstd, gamma, dgamma, dbeta, k1, k2, k3)
    dx = torch.empty_like(x, memory_format=torch.preserve_format)
    masked = None
    if want_masked:
        masked = torch.empty_like(x, memory_format=torch.preserve_format) if relu else dout
    check(lib().dcnn_bn_bwd_apply(_p(dout), _p(x), _p(out), _p(k1), _p(k2), _p(k3), _p(dx),
                                  _p(masked if (want_masked and relu) else None), LL(M), I(Cc), I(relu_mode),
                                  _stream()), "bn_bwd_apply")
    return dx, masked


def bn_inference(x, running_mean, running_var, gamma, beta, eps, residual=None, relu=False):
    x = as_nhwc(x)
    Cc = x.shape[1]
    buf = torch.empty(2 * Cc, device=x.device, dtype=torch.float32)
    scale, shift = buf[:Cc], buf[Cc:]
    check(lib().dcnn_bn_eval_coeffs(_p(running_mean), _p(running_var), _p(gamma), _p(beta), F(eps), I(Cc), _p(scale),
                                    _p(shift), _stream()), "bn_eval_coeffs")
    return bn_apply(x, None, (None, scale, shift), gamma, beta, residual, relu)


# ================================================================================== group norm
def groupnorm_forward(x, groups, gamma, beta, eps):
    x = as_nhwc(x)
    N, Cc, H, W = x.shape
    y = torch.empty_like(x, memory_format=torch.preserve_format)
    st = torch.empty(2 * N * groups, device=x.device, dtype=torch.float32)
    mean, invstd = st[:N * groups], st[N * groups:]
    check(lib().dcnn_gn_forward(_p(x), _p(y), _p(gamma), _p(beta), _p(mean), _p(invstd), I(N), I(H * W), I(Cc),
                                I(groups), F(eps), _stream()), "gn_forward")
    return y, mean, invstd


def groupnorm_backward(dy, x, mean, invstd, groups, gamma, dgamma, dbeta):
    dy = as_nhwc(dy)
    N, Cc, H, W = x.shape
    dx = torch.empty_like(x, memory_format=torch.preserve_format)
    check(lib().dcnn_gn_backward(_p(dy), _p(x), _p(mean), _p(invstd), _p(gamma), _p(dgamma), _p(dbeta), _p(dx), I(N),
                                 I(H * W), I(Cc), I(groups), _stream()), "gn_backward")
    return dx


# ================================================================================== pooling
def _pool_out(H, W, k, stride, pad):
    return (H + 2 * pad[0] - k[0]) // stride[0] + 1, (W + 2 * pad[1] - k[1]) // stride[1] + 1


def maxpool_forward(x, k, stride, pad):
    x = as_nhwc(x)
    N, Cc, H, W = x.shape
    OH, OW = _pool_out(H, W, k, stride, pad)
    y = _empty_nhwc(N, Cc, OH, OW, x.device)
    idx = torch.empty((N, OH, OW, Cc), device=x.device, dtype=torch.uint8)
    check(lib().dcnn_maxpool_fwd(_p(x), _p(y), _p(idx), I(N), I(H), I(W), I(Cc), I(OH), I(OW), I(k[0]), I(k[1]),
                                 I(stride[0]), I(stride[1]), I(pad[0]), I(pad[1]), _stream()), "maxpool_fwd")
    return y, idx


def maxpool_backward(dy, idx, x_shape, k, stride, pad):
    dy = as_nhwc(dy)
    N, Cc, H, W = x_shape
    OH, OW = dy.shape[2:]
    dx = _empty_nhwc(N, Cc, H, W, dy.device)
    check(lib().dcnn_maxpool_bwd(_p(dy), _p(idx), _p(dx), I(N), I(H), I(W), I(Cc), I(OH), I(OW), I(k[0]), I(k[1]),
                                 I(stride[0]), I(stride[1]), I(pad[0]), I(pad[1]), _stream()), "maxpool_bwd")
    return dx


def avgpool_forward(x, k, stride, pad):
    x = as_nhwc(x)
    N, Cc, H, W = x.shape
    OH, OW = _pool_out(H, W, k, stride, pad)
    y = _empty_nhwc(N, Cc, OH, OW, x.device)
    check(lib().dcnn_avgpool_fwd(_p(x), _p(y), I(N), I(H), I(W), I(Cc), I(OH), I(OW), I(k[0]), I(k[1]), I(stride[0]),
                                 I(stride[1]), I(pad[0]), I(pad[1]), _stream()), "avgpool_fwd")
    return y


def avgpool_backward(dy, x_shape, k, stride, pad):
    dy = as_nhwc(dy)
    N, Cc, H, W = x_shape
    OH, OW = dy.shape[2:]
    dx = _empty_nhwc(N, Cc, H, W, dy.device)
    check(lib().dcnn_avgpool_bwd(_p(dy), _p(dx), I(N), I(H), I(W), I(Cc), I(OH), I(OW), I(k[0]), I(k[1]),
                                 I(stride[0]), I(stride[1]), I(pad[0]), I(pad[1]), _stream()), "avgpool_bwd")
    return dx


# ================================================================================== activations
def _dense(x):
    """Make the storage dense, preserving channels-last when that is what it already is."""
    if x.dim() == 4 and x.is_contiguous(memory_format=CL):
        return x
    return x.contiguous()


def activation_forward(x, kind, alpha=0.0):
    if kind == "softmax":
        x = as_nhwc(x) if x.dim() == 4 else x.contiguous()
        Cc = x.shape[1]
        y = torch.empty_like(x, memory_format=torch.preserve_format)
        check(lib().dcnn_softmax_fwd(_p(x), _p(y), LL(x.numel() // Cc), I(Cc), _stream()), "softmax_fwd")
        return y
    x = _dense(x)
    y = torch.empty_like(x, memory_format=torch.preserve_format)
    check(lib().dcnn_act_fwd(I(_ACT[kind]), _p(x), _p(y), LL(x.numel()), F(alpha), _stream()), "act_fwd")
    return y


def activation_backward(x, dy, kind, alpha=0.0):
    if kind == "softmax":
        x = as_nhwc(x) if x.dim() == 4 else x.contiguous()
        dy = as_nhwc(dy) if dy.dim() == 4 else dy.contiguous()
        Cc = x.shape[1]
        dx = torch.empty_like(x, memory_format=torch.preserve_format)
        check(lib().dcnn_softmax_bwd(_p(x), _p(dy), _p(dx), LL(x.numel() // Cc), I(Cc), _stream()), "softmax_bwd")
        return dx
    x = _dense(x)
    dy = as_nhwc(dy) if (x.dim() == 4 and x.is_contiguous(memory_format=CL)) else dy.contiguous()
    dx = torch.empty_like(x, memory_format=torch.preserve_format)
    check(lib().dcnn_act_bwd(I(_ACT[kind]), _p(x), _p(dy), _p(dx), LL(x.numel()), F(alpha), _stream()), "act_bwd")
    return dx


def add(a, b):
    a = _dense(a)
    b = as_nhwc(b) if (a.dim() == 4 and a.is_contiguous(memory_format=CL)) else b.contiguous()
    y = torch.empty_like(a, memory_format=torch.preserve_format)
    check(lib().dcnn_add(_p(a), _p(b), _p(y), LL(a.numel()), _stream()), "add")
    return y


def dropout_forward(x, rate, seed, offset):
    x = _dense(x)
    y = torch.empty_like(x, memory_format=torch.preserve_format)
    check(lib().dcnn_dropout(_p(x), _p(y), LL(x.numel()), F(rate), ULL(seed), ULL(offset), _stream()), "dropout")
    return y


def dropout_backward(dy, rate, seed, offset):
    return dropout_forward(dy, rate, seed, offset)


# ================================================================================== layout
def flatten_nchw(x):
    N = x.shape[0]
    if x.dim() != 4 or x.shape[2] * x.shape[3] == 1 or x.shape[1] == 1:
        return x.reshape(N, -1, 1, 1)
    x = as_nhwc(x)
    _, Cc, H, W = x.shape
    out = torch.empty((N, Cc * H * W, 1, 1), device=x.device, dtype=torch.float32)
    check(lib().dcnn_transpose_batched(_p(x), _p(out), I(N), I(H * W), I(Cc), _stream()), "flatten transpose")
    return out


def unflatten_nchw(g, shape):
    N, Cc, H, W = shape
    if H * W == 1 or Cc == 1:
        return as_nhwc(g.reshape(shape))
    g = g.contiguous()
    out = _empty_nhwc(N, Cc, H, W, g.device)
    check(lib().dcnn_transpose_batched(_p(g), _p(out), I(N), I(Cc), I(H * W), _stream()), "unflatten transpose")
    return out


# ================================================================================== loss
def loss_forward_backward(kind, pred, target, param=0.0, want_grad=True):
    B = pred.shape[0]
    p = pred.reshape(B, -1).contiguous()
    t = target.reshape(B, -1).to(p.device).contiguous()
    n = p.shape[1]
    grad = torch.empty_like(p) if want_grad else None
    if kind in ("softmax_crossentropy", "logsoftmax_crossentropy", "crossentropy"):
        out = torch.empty(2, device=p.device, dtype=torch.float32)
        mode = 1 if kind == "crossentropy" else 0
        eps = (param if param > 0 else 1e-15) if mode == 1 else 0.0
        check(lib().dcnn_softmax_ce(_p(p), _p(t), _p(grad), _p(out), I(B), I(n), I(mode), F(eps), _stream()),
              "softmax_ce")
        loss, correct = out[0], out[1]
    else:
        code = {"mse": 0, "mae": 1, "huber": 2}[kind]
        out = torch.empty(1, device=p.device, dtype=torch.float32)
        check(lib().dcnn_regression_loss(_p(p), _p(t), _p(grad), _p(out), LL(p.numel()), I(code), F(param),
                                         _stream()), "regression_loss")
        loss, correct = out[0], None
    if grad is not None:
        grad = grad.reshape(B, n, 1, 1) if pred.dim() == 4 else grad.reshape(pred.shape)
    return loss, grad, correct


# ================================================================================== optimizers
_MT_TABLES = {}


def _mt_table(params, grads, s1, s2):
    key = tuple(p.data_ptr() for p in params) + tuple(g.data_ptr() for g in grads) + \
        tuple(s.data_ptr() for s in (s1 or [])) + tuple(s.data_ptr() for s in (s2 or []))
    tb = _MT_TABLES.get(key)
    if tb is None:
        dev = params[0].device
        chunk = lib().dcnn_mt_chunk()
        numel = [p.numel() for p in params]
        bt, bo = [], []
        for i, n in enumerate(numel):
            for off in range(0, n, chunk):
                bt.append(i)
                bo.append(off)

        def ptrs(ts):
            if not ts:
                return torch.zeros(len(params), dtype=torch.int64, device=dev)
            return torch.tensor([t.data_ptr() for t in ts], dtype=torch.int64, device=dev)
        for t in list(params) + list(grads) + list(s1 or []) + list(s2 or []):
            assert t.is_contiguous() or t.is_contiguous(memory_format=CL), "optimizer tensors must be dense"
        tb = (ptrs(params), ptrs(grads), ptrs(s1), ptrs(s2),
              torch.tensor(numel, dtype=torch.int64, device=dev),
              torch.tensor(bt, dtype=torch.int32, device=dev),
              torch.tensor(bo, dtype=torch.int64, device=dev), len(bt))
        if len(_MT_TABLES) > 64:
            _MT_TABLES.clear()
        _MT_TABLES[key] = tb
    return tb


def sgd_update(params, grads, velocities, lr, momentum, zero_grads=False, hyper=None):
    mode = 1 if momentum > 0 else 0
    tb = _mt_table(params, grads, velocities if mode else None, None)
    check(lib().dcnn_multi_tensor_update(I(mode), _p(tb[0]), _p(tb[1]), _p(tb[2]), _p(tb[3]), _p(tb[4]), _p(tb[5]),
                                         _p(tb[6]), I(tb[7]), F(lr), F(momentum), F(0), F(0), F(1), F(1), F(0),
                                         I(1 if zero_grads else 0), _p(hyper), _stream()), "sgd_update")
    weights_changed()


def adam_update(params, grads, ms, vs, lr, beta1, beta2, eps, bc1, bc2, weight_decay, decoupled, zero_grads=False,
                hyper=None):
    tb = _mt_table(params, grads, ms, vs)
    check(lib().dcnn_multi_tensor_update(I(3 if decoupled else 2), _p(tb[0]), _p(tb[1]), _p(tb[2]), _p(tb[3]),
                                         _p(tb[4]), _p(tb[5]), _p(tb[6]), I(tb[7]), F(lr), F(beta1), F(beta2), F(eps),
                                         F(bc1), F(bc2), F(weight_decay), I(1 if zero_grads else 0), _p(hyper),
                                         _stream()), "adam_update")
    weights_changed()
