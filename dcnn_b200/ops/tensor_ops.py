"""Tensor utilities with the reference's names (ref include/tensor/cpu/tensor_ops.hpp:22-700,
src/tensor/cuda/tensor_kernels.cu:18-308, include/ops/ops.hpp:18-945, include/math/gemm.hpp:18-36).

On the training hot path none of these run: im2col/col2im and the NCHW<->CNHW transposes are replaced by TMA-fed
implicit GEMM, padding by TMA out-of-bounds fill, residual adds / ReLU by conv and BN epilogues.  They are provided
(as thin torch expressions, device-agnostic) so that code written against the reference's tensor API has a home;
``gemm`` routes to the tcgen05 kernels when the operands qualify."""
from __future__ import annotations

from typing import List

import torch
import torch.nn.functional as F

from . import backend_for


# ---- elementwise primitives (tnn::ops::*) -------------------------------------------------------------------------
def add(a, b):
    return backend_for(a).add(a, b) if a.is_cuda and a.dtype == torch.float32 and a.shape == b.shape else a + b


def sub(a, b):
    return a - b


def mul(a, b):
    return a * b


def div(a, b):
    return a / b


def fmadd(a, b, c):
    return torch.addcmul(c, a, b)


def fmsub(a, b, c):
    return a * b - c


def fnmadd(a, b, c):
    return c - a * b


def axpy(alpha, x, y):
    """y += alpha * x (in place)."""
    return y.add_(x, alpha=alpha)


def copy(src, dst, n=None, src_offset=0, dst_offset=0):
    n = src.numel() - src_offset if n is None else n
    dst.view(-1)[dst_offset:dst_offset + n].copy_(src.reshape(-1)[src_offset:src_offset + n])
    return dst


def zero(t):
    return t.zero_()


def sum(t):  # noqa: A001 - reference name
    return t.sum()


def dot_product(a, b):
    return torch.dot(a.reshape(-1), b.reshape(-1))


def sum_squared_diff(t, mean):
    return ((t - mean) ** 2).sum()


def fill_random_uniform(t, lo, hi, seed):
    g = torch.Generator(device=t.device).manual_seed(int(seed))
    return t.uniform_(lo, hi, generator=g)


def fill_random_normal(t, mean, std, seed):
    g = torch.Generator(device=t.device).manual_seed(int(seed))
    return t.normal_(mean, std, generator=g)


def transpose_2d(t):
    return t.t().contiguous()


def nchw_to_cnhw(t):
    return t.permute(1, 0, 2, 3).contiguous()


def cnhw_to_nchw(t):
    return t.permute(1, 0, 2, 3).contiguous()


# ---- tensor ops ------------------------------------------------------------------------------------------------------
def im2col(x, kernel_h, kernel_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0):
    """-> [C*kh*kw, N*OH*OW] (the reference's column layout)."""
    cols = F.unfold(x, (kernel_h, kernel_w), padding=(pad_h, pad_w), stride=(stride_h, stride_w))  # [N, C*k, L]
    return cols.permute(1, 0, 2).reshape(cols.shape[1], -1)


def col2im(cols, n, c, h, w, kernel_h, kernel_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0):
    L = cols.shape[1] // n
    u = cols.reshape(cols.shape[0], n, L).permute(1, 0, 2)
    return F.fold(u, (h, w), (kernel_h, kernel_w), padding=(pad_h, pad_w), stride=(stride_h, stride_w))


def pad(x, pad_h, pad_w, value=0.0):
    return F.pad(x, (pad_w, pad_w, pad_h, pad_h), value=value)


def unpad(x, pad_h, pad_w):
    return x[:, :, pad_h:x.shape[2] - pad_h, pad_w:x.shape[3] - pad_w].contiguous()


def crop(x, start_h, start_w, end_h, end_w):
    """Inclusive end indices, like the reference."""
    return x[:, :, start_h:end_h + 1, start_w:end_w + 1].contiguous()


def slice_batch(x, start, end):
    return x[start:end].contiguous()


def slice_channels(x, start, end):
    """Inclusive end channel, like the reference."""
    return x[:, start:end + 1].contiguous()


def split(x, num_splits: int) -> List[torch.Tensor]:
    """B/M each, the last split takes the remainder (tensor_ops.hpp:675-689)."""
    from ..train import split_microbatches
    if num_splits <= 0 or num_splits > x.shape[0]:
        raise ValueError("Invalid number of splits")
    return [t.contiguous() for t in split_microbatches(x, num_splits)]


def apply_softmax(x):
    """Channel softmax, in place."""
    x.copy_(torch.softmax(x, dim=1))
    return x


# ---- scalar / comparison primitives and tensor helpers (tnn::ops::*, Tensor<T> members) -----------------------------
def add_scalar(a, scalar):
    return a + scalar


def sub_scalar(a, scalar):
    return a - scalar


def mul_scalar(a, scalar):
    return a * scalar


def div_scalar(a, scalar):
    if scalar == 0:
        raise ZeroDivisionError("Division by zero")
    return a / scalar


def set_scalar(t, scalar):
    t.fill_(scalar)
    return t


def mul_add_scalar(a, mul, add_):
    """a * mul + add"""
    return torch.add(a * mul, add_)


def sub_mul_scalar(a, sub_, mul):
    """(a - sub) * mul"""
    return (a - sub_) * mul


def scalar_max(a, scalar):
    return torch.clamp_min(a, scalar)


def clamp(a, min_val, max_val):
    return torch.clamp(a, min_val, max_val)


def equal(a, b):
    """1.0 where a == b else 0.0 (the reference's comparison ops produce masks of the value type)."""
    return (a == b).to(a.dtype)


def greater(a, b):
    return (a > b).to(a.dtype)


def rsqrt(a):
    return torch.rsqrt(a)


def norm_squared(t) -> float:
    return float((t.double() * t.double()).sum())


def mean(t) -> float:
    return float(t.double().mean())


def variance(t) -> float:
    """Biased variance of all elements (Tensor::variance)."""
    return float(t.double().var(unbiased=False))


def same_shape(a, b) -> bool:
    return tuple(a.shape) == tuple(b.shape)


def shape_str(t) -> str:
    return "(" + ", ".join(str(int(d)) for d in t.shape) + ")"


def check_nan_and_inf(t, name: str = "tensor") -> bool:
    """True (and a diagnostic line) if the tensor holds a NaN or an infinity (Tensor::check_nan_and_inf)."""
    bad = ~torch.isfinite(t)
    if bool(bad.any()):
        idx = torch.nonzero(bad.reshape(-1))[0].item()
        print(f"{name}: non-finite value {float(t.reshape(-1)[idx])} at flat index {idx} "
              f"({int(bad.sum())} of {t.numel()} elements)")
        return True
    return False


def copy_batch(src, src_batch: int, dst, dst_batch: int):
    """dst[dst_batch] = src[src_batch] (Tensor::copy_batch)."""
    if tuple(src.shape[1:]) != tuple(dst.shape[1:]):
        raise ValueError("copy_batch: per-sample shapes differ")
    dst[dst_batch].copy_(src[src_batch])
    return dst


def print_data(t, max_elements: int = 64):
    flat = t.detach().reshape(-1)[:max_elements].tolist()
    more = " ..." if t.numel() > max_elements else ""
    print(f"Tensor{shape_str(t)} on {t.device}: " + " ".join(f"{v:.6g}" for v in flat) + more)


# ---- GEMM ------------------------------------------------------------------------------------------------------------
def gemm(a, b, trans_a=False, trans_b=False, alpha=1.0, beta=0.0, c=None):
    """C = alpha * op(A) op(B) + beta * C for row-major 2-D tensors (ref math/gemm.hpp).  On CUDA, fp32 ``A @ B^T`` with K a
    multiple of 32 runs on the tcgen05 Dense kernel; every other case is a plain torch matmul (not a hot path)."""
    A = a.t() if trans_a else a
    if a.is_cuda and a.dtype == torch.float32 and A.shape[1] % 32 == 0 and b.dtype == torch.float32:
        Bt = b if trans_b else b.t()          # the Dense kernel wants the second operand as [N, K]
        out = backend_for(a).dense_forward(A.contiguous(), Bt.contiguous(), None)
    else:
        out = A @ (b.t() if trans_b else b)
    if alpha != 1.0:
        out = out * alpha
    if c is not None and beta != 0.0:
        out = out + beta * c
    if c is not None:
        c.copy_(out)
        return c
    return out
