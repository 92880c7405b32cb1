"""Native CPU math (libdcnn_host.so, csrc/host/cpu_math.cpp): blocked AVX2/FMA sgemm/dgemm, vectorised elementwise
primitives, reductions, dot product, im2col/col2im -- the counterpart of the reference's hand-written CPU backend
(ref src/math/cpu/sgemm.cpp, dgemm.cpp; src/ops/cpu/skernels.cpp, dkernels.cpp; src/utils/simd_dot_product.S;
include/threading/thread_handler.hpp; include/tensor/cpu/tensor_ops.hpp:22-330).

``DCNN_CPU_BACKEND=native`` routes the CPU path's convolution (im2col + GEMM, as the reference's CPU Conv2D does:
src/nn/layers_impl/cpu/conv2d_ops.cpp:19-67), Dense, BatchNorm and pooling layers (csrc/host/cpu_layers.cpp; ref
batchnorm_ops.cpp, maxpool_ops.cpp, avgpool_ops.cpp) through these kernels instead of ATen; the default stays
ATen (oneDNN/MKL are faster on most hosts; the native path exists for parity and for hosts without them)."""
from __future__ import annotations

import ctypes as C
import os

import torch

from .._hostlib import lib as _hostlib

BIN = {"add": 0, "sub": 1, "mul": 2, "div": 3, "max": 4, "min": 5}
UN = {"neg": 0, "abs": 1, "sqrt": 2, "rsqrt": 3, "rcp": 4, "exp": 5, "log": 6, "square": 7, "relu": 8}
_SETUP = False


def _lib():
    global _SETUP
    l = _hostlib()
    if not _SETUP:
        l.dcnn_cpu_sdot.restype = C.c_float
        l.dcnn_cpu_ddot.restype = C.c_double
        l.dcnn_cpu_ssum.restype = C.c_double
        l.dcnn_cpu_ssum_squared_diff.restype = C.c_double
        l.dcnn_cpu_get_threads.restype = C.c_int
        _SETUP = True
    return l


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _chk(*ts):
    for t in ts:
        if t.is_cuda or not t.is_contiguous() or t.dtype not in (torch.float32, torch.float64):
            raise ValueError("cpu_native ops take contiguous float32/float64 CPU tensors")
    if len({t.dtype for t in ts}) != 1:
        raise ValueError("mixed dtypes")
    return ts[0].dtype == torch.float64


def set_num_threads(n: int):
    """ref ThreadWrapper (include/threading/thread_wrapper.hpp:18-47): bound the worker count of the native kernels."""
    _lib().dcnn_cpu_set_threads(C.c_int(int(n)))


def get_num_threads() -> int:
    return int(_lib().dcnn_cpu_get_threads())


def gemm(a, b, trans_a=False, trans_b=False, alpha=1.0, beta=0.0, c=None):
    """C = alpha * op(A) @ op(B) + beta * C, row-major 2-D float32/float64 (ref include/math/cpu/gemm.hpp:14-32)."""
    dbl = _chk(a, b)
    M, K = (a.shape[1], a.shape[0]) if trans_a else a.shape
    K2, N = (b.shape[1], b.shape[0]) if trans_b else b.shape
    if K != K2:
        raise ValueError(f"gemm: inner dimensions differ ({K} vs {K2})")
    if c is None:
        c = torch.empty(M, N, dtype=a.dtype)
        beta = 0.0
    else:
        _chk(a, c)
        if tuple(c.shape) != (M, N):
            raise ValueError("gemm: bad output shape")
    fn = _lib().dcnn_cpu_dgemm if dbl else _lib().dcnn_cpu_sgemm
    sc = C.c_double if dbl else C.c_float
    fn(C.c_int(int(trans_a)), C.c_int(int(trans_b)), C.c_int(M), C.c_int(N), C.c_int(K), sc(alpha), _ptr(a),
       C.c_int(a.shape[1]), _ptr(b), C.c_int(b.shape[1]), sc(beta), _ptr(c), C.c_int(N))
    return c


def dot(a, b) -> float:
    dbl = _chk(a, b)
    n = C.c_longlong(a.numel())
    return float((_lib().dcnn_cpu_ddot if dbl else _lib().dcnn_cpu_sdot)(_ptr(a), _ptr(b), n))


def sum(a) -> float:  # noqa: A001 - reference name
    if _chk(a):
        return float(a.sum())
    return float(_lib().dcnn_cpu_ssum(_ptr(a), C.c_longlong(a.numel())))


def sum_squared_diff(a, mean: float) -> float:
    if _chk(a):
        return float(((a - mean) ** 2).sum())
    return float(_lib().dcnn_cpu_ssum_squared_diff(_ptr(a), C.c_float(mean), C.c_longlong(a.numel())))


def binary(op: str, a, b, out=None):
    """add / sub / mul / div / max / min; ``b`` may be a Python scalar."""
    scalar = not torch.is_tensor(b)
    dbl = _chk(a) if scalar else _chk(a, b)
    out = torch.empty_like(a) if out is None else out
    n = C.c_longlong(a.numel())
    l = _lib()
    if scalar:
        (l.dcnn_cpu_dbinary_scalar if dbl else l.dcnn_cpu_sbinary_scalar)(
            C.c_int(BIN[op]), _ptr(a), (C.c_double if dbl else C.c_float)(b), _ptr(out), n)
    else:
        if a.shape != b.shape:
            raise ValueError("binary: shapes differ")
        (l.dcnn_cpu_dbinary if dbl else l.dcnn_cpu_sbinary)(C.c_int(BIN[op]), _ptr(a), _ptr(b), _ptr(out), n)
    return out


def unary(op: str, a, out=None):
    dbl = _chk(a)
    out = torch.empty_like(a) if out is None else out
    (_lib().dcnn_cpu_dunary if dbl else _lib().dcnn_cpu_sunary)(C.c_int(UN[op]), _ptr(a), _ptr(out),
                                                                C.c_longlong(a.numel()))
    return out


def fma(a, b, c, mode="fmadd", out=None):
    """fmadd: a*b+c, fmsub: a*b-c, fnmadd: c-a*b (float32)."""
    _chk(a, b, c)
    out = torch.empty_like(a) if out is None else out
    _lib().dcnn_cpu_sfma(C.c_int({"fmadd": 0, "fmsub": 1, "fnmadd": 2}[mode]), _ptr(a), _ptr(b), _ptr(c), _ptr(out),
                         C.c_longlong(a.numel()))
    return out


def axpy(alpha, x, y):
    _chk(x, y)
    _lib().dcnn_cpu_saxpy(C.c_float(alpha), _ptr(x), _ptr(y), C.c_longlong(x.numel()))
    return y


def fill(y, value):
    _chk(y)
    _lib().dcnn_cpu_sset(_ptr(y), C.c_float(value), C.c_longlong(y.numel()))
    return y


def transpose_2d(a):
    _chk(a)
    out = torch.empty(a.shape[1], a.shape[0], dtype=a.dtype)
    _lib().dcnn_cpu_stranspose(_ptr(a), _ptr(out), C.c_int(a.shape[0]), C.c_int(a.shape[1]))
    return out


def _out_hw(H, W, kh, kw, stride, pad):
    return (H + 2 * pad[0] - kh) // stride[0] + 1, (W + 2 * pad[1] - kw) // stride[1] + 1


def im2col(x, kh, kw, stride=(1, 1), pad=(0, 0)):
    """[N,C,H,W] -> [C*kh*kw, N*OH*OW]"""
    x = x.contiguous()
    _chk(x)
    N, Cc, H, W = x.shape
    OH, OW = _out_hw(H, W, kh, kw, stride, pad)
    col = torch.empty(Cc * kh * kw, N * OH * OW, dtype=torch.float32)
    _lib().dcnn_cpu_im2col(_ptr(x), _ptr(col), *(C.c_int(v) for v in (N, Cc, H, W, kh, kw, stride[0], stride[1],
                                                                       pad[0], pad[1], OH, OW)))
    return col


def col2im(col, x_shape, kh, kw, stride=(1, 1), pad=(0, 0)):
    col = col.contiguous()
    _chk(col)
    N, Cc, H, W = x_shape
    OH, OW = _out_hw(H, W, kh, kw, stride, pad)
    x = torch.empty(N, Cc, H, W, dtype=torch.float32)
    _lib().dcnn_cpu_col2im(_ptr(col), _ptr(x), *(C.c_int(v) for v in (N, Cc, H, W, kh, kw, stride[0], stride[1],
                                                                       pad[0], pad[1], OH, OW)))
    return x


# ---------------------------------------------------------------------------------------------- layer ops on the native GEMM
def enabled() -> bool:
    return os.environ.get("DCNN_CPU_BACKEND", "aten").lower() == "native"


def _conv_geometry(x_shape, w_shape, stride, pad):
    N, Cc, H, W = x_shape
    OC, _, kh, kw = w_shape
    OH, OW = _out_hw(H, W, kh, kw, stride, pad)
    return [C.c_int(int(v)) for v in (N, Cc, H, W, OC, kh, kw, stride[0], stride[1], pad[0], pad[1], OH, OW)], OH, OW


def conv2d_forward(x, w, bias, stride, pad):
    """Blocked im2col + GEMM, the output written straight into NCHW (csrc/host/cpu_conv.cpp)."""
    x, w = x.float().contiguous(), w.contiguous()
    geo, OH, OW = _conv_geometry(x.shape, w.shape, stride, pad)
    y = torch.empty(x.shape[0], w.shape[0], OH, OW)
    b = bias.contiguous() if bias is not None else None
    _lib().dcnn_cpu_conv2d_fwd(_ptr(x), _ptr(w), _ptr(b) if b is not None else C.c_void_p(0), _ptr(y), *geo)
    return y


def conv2d_backward_data(dy, w, x_shape, stride, pad):
    dy, w = dy.float().contiguous(), w.contiguous()
    geo, _, _ = _conv_geometry(x_shape, w.shape, stride, pad)
    dx = torch.empty(*x_shape)
    _lib().dcnn_cpu_conv2d_bwd(C.c_void_p(0), _ptr(w), _ptr(dy), C.c_void_p(0), _ptr(dx), *geo)
    return dx


def conv2d_backward_weight(dy, x, dw, stride, pad):
    """dw += dy (*) x, accumulated in place when ``dw`` is contiguous."""
    dy, x = dy.float().contiguous(), x.float().contiguous()
    geo, _, _ = _conv_geometry(x.shape, dw.shape, stride, pad)
    acc = dw if dw.is_contiguous() else torch.zeros(dw.shape)
    _lib().dcnn_cpu_conv2d_bwd(_ptr(x), C.c_void_p(0), _ptr(dy), _ptr(acc), C.c_void_p(0), *geo)
    if acc is not dw:
        dw.add_(acc)


def dense_forward(x2d, w2d, bias):
    y = gemm(x2d.contiguous(), w2d.contiguous(), trans_b=True)
    if bias is not None:
        y += bias.view(1, -1)
    return y


def dense_backward_data(dy2d, w2d):
    return gemm(dy2d.contiguous(), w2d.contiguous())


def dense_backward_weight(dy2d, x2d, dw2d):
    if dw2d.is_contiguous():
        gemm(dy2d.contiguous(), x2d.contiguous(), trans_a=True, beta=1.0, c=dw2d)
    else:
        dw2d.add_(gemm(dy2d.contiguous(), x2d.contiguous(), trans_a=True))


# ---------------------------------------------------------------------------------------------- native layer kernels
def _ci(*v):
    return [C.c_int(int(a)) for a in v]


def bn_stats(x):
    """-> (mean[C], biased var[C]) over (N, H, W)."""
    x = x.contiguous()
    n, c = x.shape[:2]
    hw = x.numel() // (n * c)
    mean, var = torch.empty(c), torch.empty(c)
    _lib().dcnn_cpu_bn_stats(_ptr(x), *_ci(n, c, hw), _ptr(mean), _ptr(var))
    return mean, var


def bn_apply(x, scale, shift, residual=None, relu=False):
    x = x.contiguous()
    n, c = x.shape[:2]
    hw = x.numel() // (n * c)
    y = torch.empty_like(x)
    r = residual.contiguous() if residual is not None else None
    _lib().dcnn_cpu_bn_apply(_ptr(x), _ptr(scale.contiguous()), _ptr(shift.contiguous()),
                             _ptr(r) if r is not None else C.c_void_p(0), _ptr(y), *_ci(n, c, hw, 1 if relu else 0))
    return y


def bn_backward_sums(dout, x, out, mean, invstd, relu):
    dout, x = dout.contiguous(), x.contiguous()
    n, c = x.shape[:2]
    hw = x.numel() // (n * c)
    a, b = torch.empty(c, dtype=torch.float64), torch.empty(c, dtype=torch.float64)
    o = out.contiguous() if (relu and out is not None) else None
    _lib().dcnn_cpu_bn_bwd_sums(_ptr(dout), _ptr(x), _ptr(o) if o is not None else C.c_void_p(0), _ptr(mean.contiguous()),
                                _ptr(invstd.contiguous()), *_ci(n, c, hw, 1 if o is not None else 0), _ptr(a), _ptr(b))
    return a, b


def bn_backward_dx(dout, x, out, mean, invstd, scale, sum_dy, sum_dy_xhat, count, relu, want_masked):
    dout, x = dout.contiguous(), x.contiguous()
    n, c = x.shape[:2]
    hw = x.numel() // (n * c)
    dx = torch.empty_like(x)
    o = out.contiguous() if (relu and out is not None) else None
    masked = torch.empty_like(x) if want_masked else None
    _lib().dcnn_cpu_bn_bwd_dx(_ptr(dout), _ptr(x), _ptr(o) if o is not None else C.c_void_p(0), _ptr(mean.contiguous()),
                              _ptr(invstd.contiguous()), _ptr(scale.contiguous()), _ptr(sum_dy.contiguous()),
                              _ptr(sum_dy_xhat.contiguous()), C.c_double(float(count)),
                              *_ci(n, c, hw, 1 if o is not None else 0), _ptr(dx),
                              _ptr(masked) if masked is not None else C.c_void_p(0))
    return dx, masked


def maxpool_forward(x, k, stride, pad):
    x = x.contiguous()
    n, c, h, w = x.shape
    oh, ow = _out_hw(h, w, k[0], k[1], stride, pad)
    y = torch.empty(n, c, oh, ow)
    idx = torch.empty(n, c, oh, ow, dtype=torch.int64)
    _lib().dcnn_cpu_maxpool_fwd(_ptr(x), _ptr(y), _ptr(idx),
                                *_ci(n, c, h, w, oh, ow, k[0], k[1], stride[0], stride[1], pad[0], pad[1]))
    return y, idx


def maxpool_backward(dy, idx, x_shape):
    dy, idx = dy.contiguous(), idx.contiguous()
    n, c, h, w = x_shape
    dx = torch.empty(n, c, h, w)
    _lib().dcnn_cpu_maxpool_bwd(_ptr(dy), _ptr(idx), _ptr(dx), *_ci(n, c, h * w, dy.shape[2] * dy.shape[3]))
    return dx


def avgpool_forward(x, k, stride, pad):
    x = x.contiguous()
    n, c, h, w = x.shape
    oh, ow = _out_hw(h, w, k[0], k[1], stride, pad)
    y = torch.empty(n, c, oh, ow)
    _lib().dcnn_cpu_avgpool_fwd(_ptr(x), _ptr(y), *_ci(n, c, h, w, oh, ow, k[0], k[1], stride[0], stride[1], pad[0],
                                                     pad[1]))
    return y


def avgpool_backward(dy, x_shape, k, stride, pad):
    dy = dy.contiguous()
    n, c, h, w = x_shape
    dx = torch.empty(n, c, h, w)
    _lib().dcnn_cpu_avgpool_bwd(_ptr(dy), _ptr(dx), *_ci(n, c, h, w, dy.shape[2], dy.shape[3], k[0], k[1], stride[0],
                                                       stride[1], pad[0], pad[1]))
    return dx


def sgd_update(params, grads, velocities, lr, momentum, zero_grads=False):
    for i, (p, g) in enumerate(zip(params, grads)):
        v = velocities[i] if (velocities is not None and momentum > 0) else None
        _lib().dcnn_cpu_sgd_update(_ptr(p), _ptr(g), _ptr(v) if v is not None else C.c_void_p(0),
                                   C.c_longlong(p.numel()), C.c_float(lr), C.c_float(momentum),
                                   C.c_int(1 if zero_grads else 0))


def adam_update(params, grads, ms, vs, lr, beta1, beta2, eps, bc1, bc2, weight_decay, decoupled, zero_grads=False):
    for p, g, m, v in zip(params, grads, ms, vs):
        _lib().dcnn_cpu_adam_update(_ptr(p), _ptr(g), _ptr(m), _ptr(v), C.c_longlong(p.numel()), C.c_float(lr),
                                    C.c_float(beta1), C.c_float(beta2), C.c_float(eps), C.c_float(bc1), C.c_float(bc2),
                                    C.c_float(weight_decay), C.c_int(1 if decoupled else 0),
                                    C.c_int(1 if zero_grads else 0))


def softmax_ce(pred2d, target2d, want_grad=True):
    """-> (loss tensor, grad [B, C] | None, correct-count tensor)."""
    p, t = pred2d.contiguous(), target2d.contiguous()
    b, c = p.shape
    grad = torch.empty_like(p) if want_grad else None
    out = torch.empty(2, dtype=torch.float64)
    _lib().dcnn_cpu_softmax_ce(_ptr(p), _ptr(t), _ptr(grad) if grad is not None else C.c_void_p(0), _ptr(out),
                               C.c_int(b), C.c_int(c))
    return out[0].to(torch.float32), grad, out[1].to(torch.int64)
