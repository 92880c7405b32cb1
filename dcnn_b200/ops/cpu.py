"""CPU (ATen, FP32) implementations of the functional op set.

This backend is the numerics oracle for every CUDA kernel test and the execution path of the
CPU-only plumbing config (``mnist_cnn_trainer`` with ``DEVICE_TYPE=CPU``).  Semantics follow the
reference's CPU kernels (e.g. /root/reference/src/nn/layers_impl/cpu/batchnorm_ops.cpp:16-110,
conv2d_ops.cpp:19-67, maxpool_ops / avgpool_ops, loss_ops.cpp) but are written against ATen.

All activation tensors are logical ``[N, C, H, W]``.
"""
from __future__ import annotations


import torch
import torch.nn.functional as F

NAME = "cpu"


# ----------------------------------------------------------------------------------------------
# convolution / dense
# ----------------------------------------------------------------------------------------------
class CpuStats:
    """Per-channel batch statistics (mean, biased var, count)."""

    __slots__ = ("mean", "var", "count")

    def __init__(self, mean, var, count):
        self.mean, self.var, self.count = mean, var, count


def _native():
    """The hand-written AVX2 GEMM path (ops/cpu_native.py) when DCNN_CPU_BACKEND=native, else None (ATen)."""
    from . import cpu_native
    return cpu_native if cpu_native.enabled() else None


def conv2d_forward(x, w, bias, stride, pad, want_stats=False, relu=False):
    nat = _native()
    if nat is not None and x.dtype == torch.float32:
        y = nat.conv2d_forward(x, w, bias, stride, pad)
    else:
        y = F.conv2d(x, w, None if bias is None else bias.reshape(-1), stride=stride, padding=pad)
    if relu:
        y = torch.relu_(y)
    return y, (bn_stats(y) if want_stats else None)


def conv2d_backward_data(dy, w, x_shape, stride, pad, out=None, accumulate=False):
    nat = _native()
    if nat is not None and dy.dtype == torch.float32:
        dx = nat.conv2d_backward_data(dy, w, tuple(x_shape), stride, pad)
    else:
        dx = torch.nn.grad.conv2d_input(list(x_shape), w, dy, stride=stride, padding=pad)
    if out is not None:
        if accumulate:
            out.add_(dx)
        else:
            out.copy_(dx)
        return out
    return dx


def conv2d_backward_weight(dy, x, dw, stride, pad, db=None):
    """dw += dL/dW  (beta=1 accumulation across microbatches, ref cudnn_conv2d_ops.cu:226); db += bias gradient."""
    if db is not None:
        bias_grad(dy, db)
    nat = _native()
    if nat is not None and dy.dtype == torch.float32:
        return nat.conv2d_backward_weight(dy, x, dw, stride, pad)
    g = torch.nn.grad.conv2d_weight(x, list(dw.shape), dy, stride=stride, padding=pad)
    dw.add_(g)


def bias_grad(dy, db):
    """db += sum of dy over every axis but channels."""
    db.add_(dy.sum(dim=(0, 2, 3)).view_as(db))


def dense_forward(x2d, w2d, bias, relu=False):
    nat = _native()
    if nat is not None and x2d.dtype == torch.float32:
        y = nat.dense_forward(x2d, w2d, bias)
    else:
        y = x2d @ w2d.t()
        if bias is not None:
            y = y + bias.view(1, -1)
    return torch.relu_(y) if relu else y


def dense_backward_data(dy2d, w2d):
    nat = _native()
    if nat is not None and dy2d.dtype == torch.float32:
        return nat.dense_backward_data(dy2d, w2d)
    return dy2d @ w2d


def dense_backward_weight(dy2d, x2d, dw2d, db=None):
    if db is not None:
        db.add_(dy2d.sum(0).view_as(db))
    nat = _native()
    if nat is not None and dy2d.dtype == torch.float32:
        return nat.dense_backward_weight(dy2d, x2d, dw2d)
    dw2d.add_(dy2d.t() @ x2d)


# ----------------------------------------------------------------------------------------------
# batch norm
# ----------------------------------------------------------------------------------------------
def bn_stats(x):
    n = x.numel() // x.shape[1]
    nat = _native()
    if nat is not None and x.dtype == torch.float32 and x.dim() == 4:
        mean, var = nat.bn_stats(x)
        return CpuStats(mean, var, n)
    mean = x.mean(dim=(0, 2, 3))
    var = x.var(dim=(0, 2, 3), unbiased=False)
    return CpuStats(mean, var, n)


def bn_finalize(stats, eps, momentum, running_mean, running_var, sync=None, gamma=None, beta=None):
    """Turn partial statistics into (mean, invstd) and update running stats.

    Normalisation uses the biased variance, the running variance gets the unbiased one
    (ref batchnorm_ops.cu:141-153)."""
    mean, var, n = stats.mean, stats.var, stats.count
    if sync is not None:
        mean, var, n = sync.merge_cpu(mean, var, n)
    invstd = torch.rsqrt(var + eps)
    if running_mean is not None:
        unbiased = var * (n / max(n - 1, 1))
        running_mean.mul_(1 - momentum).add_(mean, alpha=momentum)
        running_var.mul_(1 - momentum).add_(unbiased, alpha=momentum)
    return mean, invstd


def bn_apply(x, mean, invstd, gamma, beta, residual=None, relu=False):
    c = x.shape[1]
    scale = invstd if gamma is None else invstd * gamma.view(-1)
    shift = -mean * scale
    if beta is not None:
        shift = shift + beta.view(-1)
    nat = _native()
    if nat is not None and x.dtype == torch.float32 and x.dim() == 4:
        return nat.bn_apply(x, scale, shift, residual, relu)
    y = x * scale.view(1, c, 1, 1) + shift.view(1, c, 1, 1)
    if residual is not None:
        y = y + residual
    if relu:
        y = torch.relu(y)
    return y


def bn_backward(dout, x, out, mean, invstd, gamma, dgamma, dbeta, relu=False, want_masked=False,
                sync=None, recompute_mask=False, relu_mask=None):
    """Backward of (BN -> [+residual] -> [ReLU]).

    ``out`` is the forward output (needed for the ReLU mask).  Returns (dx, masked_dout|None).
    dgamma/dbeta are accumulated (deviation from ref batchnorm_ops.cu:218-219, which overwrites;
    see SURVEY appendix B.2)."""
    c = x.shape[1]
    nat = _native()
    if nat is not None and x.dtype == torch.float32 and x.dim() == 4:
        m = x.numel() // c
        a, b = nat.bn_backward_sums(dout, x, out, mean, invstd, relu)
        if dgamma is not None:
            dgamma.add_(b.float().view_as(dgamma))
            dbeta.add_(a.float().view_as(dbeta))
        if sync is not None:
            a32, b32, m = sync.merge_bwd_cpu(a.float(), b.float(), m)
            a, b = a32.double(), b32.double()
        scale = invstd if gamma is None else invstd * gamma.view(-1)
        return nat.bn_backward_dx(dout, x, out, mean, invstd, scale, a, b, m, relu, want_masked)
    g = dout
    if relu:
        g = dout * (out > 0).to(dout.dtype)
    xhat = (x - mean.view(1, c, 1, 1)) * invstd.view(1, c, 1, 1)
    sum_dy = g.sum(dim=(0, 2, 3))
    sum_dy_xhat = (g * xhat).sum(dim=(0, 2, 3))
    m = x.numel() // c
    if dgamma is not None:  # local contribution; the data-parallel gradient all-reduce averages the replicas
        dgamma.add_(sum_dy_xhat.view_as(dgamma))
        dbeta.add_(sum_dy.view_as(dbeta))
    if sync is not None:
        sum_dy, sum_dy_xhat, m = sync.merge_bwd_cpu(sum_dy, sum_dy_xhat, m)
    scale = invstd if gamma is None else invstd * gamma.view(-1)
    dx = (g - (sum_dy / m).view(1, c, 1, 1) - xhat * (sum_dy_xhat / m).view(1, c, 1, 1)) \
        * scale.view(1, c, 1, 1)
    return dx, (g if want_masked else None)


def bn_inference(x, running_mean, running_var, gamma, beta, eps, residual=None, relu=False):
    invstd = torch.rsqrt(running_var + eps)
    return bn_apply(x, running_mean, invstd, gamma, beta, residual, relu)


# ----------------------------------------------------------------------------------------------
# group norm
# ----------------------------------------------------------------------------------------------
def groupnorm_forward(x, groups, gamma, beta, eps):
    n, c, h, w = x.shape
    xg = x.reshape(n, groups, -1)
    mean = xg.mean(dim=2)
    var = xg.var(dim=2, unbiased=False)
    invstd = torch.rsqrt(var + eps)
    xhat = ((xg - mean.unsqueeze(2)) * invstd.unsqueeze(2)).reshape(n, c, h, w)
    y = xhat
    if gamma is not None:
        y = xhat * gamma.view(1, c, 1, 1) + beta.view(1, c, 1, 1)
    return y, mean, invstd


def groupnorm_backward(dy, x, mean, invstd, groups, gamma, dgamma, dbeta):
    n, c, h, w = x.shape
    xg = x.reshape(n, groups, -1)
    xhat = ((xg - mean.unsqueeze(2)) * invstd.unsqueeze(2)).reshape(n, c, h, w)
    if dgamma is not None:
        dgamma.add_((dy * xhat).sum(dim=(0, 2, 3)).view_as(dgamma))
        dbeta.add_(dy.sum(dim=(0, 2, 3)).view_as(dbeta))
    dxhat = dy if gamma is None else dy * gamma.view(1, c, 1, 1)
    dxg = dxhat.reshape(n, groups, -1)
    xhg = xhat.reshape(n, groups, -1)
    m = dxg.shape[2]
    s1 = dxg.sum(dim=2, keepdim=True)
    s2 = (dxg * xhg).sum(dim=2, keepdim=True)
    dx = (dxg - s1 / m - xhg * s2 / m) * invstd.unsqueeze(2)
    return dx.reshape(n, c, h, w)


# ----------------------------------------------------------------------------------------------
# pooling
# ----------------------------------------------------------------------------------------------
def maxpool_forward(x, k, stride, pad):
    nat = _native()
    if nat is not None and x.dtype == torch.float32:
        return nat.maxpool_forward(x, k, stride, pad)
    y, idx = F.max_pool2d(x, k, stride, pad, return_indices=True)
    return y, idx


def maxpool_backward(dy, idx, x_shape, k, stride, pad):
    nat = _native()
    if nat is not None and dy.dtype == torch.float32:
        return nat.maxpool_backward(dy, idx, tuple(x_shape))
    return F.max_unpool2d(dy, idx, k, stride, pad, output_size=list(x_shape[2:])) \
        if _non_overlapping(k, stride) else _maxpool_bwd_scatter(dy, idx, x_shape)


def _non_overlapping(k, stride):
    return stride[0] >= k[0] and stride[1] >= k[1]


def _maxpool_bwd_scatter(dy, idx, x_shape):
    n, c = x_shape[:2]
    dx = torch.zeros(n, c, x_shape[2] * x_shape[3], dtype=dy.dtype, device=dy.device)
    dx.scatter_add_(2, idx.reshape(n, c, -1), dy.reshape(n, c, -1))
    return dx.reshape(x_shape)


def avgpool_forward(x, k, stride, pad):
    # divisor is always pool_h*pool_w (ref avgpool_ops.cu:106,122)
    nat = _native()
    if nat is not None and x.dtype == torch.float32:
        return nat.avgpool_forward(x, k, stride, pad)
    return F.avg_pool2d(x, k, stride, pad, count_include_pad=True)


def avgpool_backward(dy, x_shape, k, stride, pad):
    nat = _native()
    if nat is not None and dy.dtype == torch.float32:
        return nat.avgpool_backward(dy, tuple(x_shape), k, stride, pad)
    n, c, h, w = x_shape
    ones = torch.ones(c, 1, k[0], k[1], dtype=dy.dtype, device=dy.device) / (k[0] * k[1])
    oh, ow = dy.shape[2:]
    nat_h = (oh - 1) * stride[0] - 2 * pad[0] + k[0]
    nat_w = (ow - 1) * stride[1] - 2 * pad[1] + k[1]
    op = (max(0, min(h - nat_h, stride[0] - 1)), max(0, min(w - nat_w, stride[1] - 1)))
    dx = F.conv_transpose2d(dy, ones, stride=stride, padding=pad, output_padding=op, groups=c)
    if dx.shape[2] != h or dx.shape[3] != w:  # rows/cols no window reaches get zero gradient
        full = torch.zeros(x_shape, dtype=dy.dtype, device=dy.device)
        hh, ww = min(h, dx.shape[2]), min(w, dx.shape[3])
        full[:, :, :hh, :ww] = dx[:, :, :hh, :ww]
        dx = full
    return dx


# ----------------------------------------------------------------------------------------------
# activations  (grads recomputed from the pre-activation input, ref activations_impl/*)
# ----------------------------------------------------------------------------------------------
def activation_forward(x, kind, alpha=0.0):
    if kind == "relu":
        return torch.relu(x)
    if kind == "leaky_relu":
        return F.leaky_relu(x, alpha)
    if kind == "elu":
        return F.elu(x, alpha)
    if kind == "sigmoid":
        return torch.sigmoid(x)
    if kind == "tanh":
        return torch.tanh(x)
    if kind == "softmax":
        return torch.softmax(x, dim=1)
    if kind == "linear":
        return x.clone()
    raise ValueError(f"unknown activation {kind}")


def activation_backward(x, dy, kind, alpha=0.0):
    if kind == "relu":
        return dy * (x > 0).to(dy.dtype)
    if kind == "leaky_relu":
        return dy * torch.where(x > 0, torch.ones_like(x), torch.full_like(x, alpha))
    if kind == "elu":
        return dy * torch.where(x > 0, torch.ones_like(x), alpha * torch.exp(x))
    if kind == "sigmoid":
        s = torch.sigmoid(x)
        return dy * s * (1 - s)
    if kind == "tanh":
        t = torch.tanh(x)
        return dy * (1 - t * t)
    if kind == "softmax":
        s = torch.softmax(x, dim=1)
        dot = (s * dy).sum(dim=1, keepdim=True)
        return s * (dy - dot)
    if kind == "linear":
        return dy.clone()
    raise ValueError(f"unknown activation {kind}")


def add(a, b):
    return a + b


# ----------------------------------------------------------------------------------------------
# dropout
# ----------------------------------------------------------------------------------------------
def dropout_forward(x, rate, seed, offset):
    gen = torch.Generator(device=x.device)
    gen.manual_seed((seed * 1000003 + offset) & 0x7FFFFFFFFFFFFFFF)
    keep = (torch.rand(x.shape, generator=gen, device=x.device) >= rate).to(x.dtype)
    scale = 1.0 / (1.0 - rate)
    return x * keep * scale


def dropout_backward(dy, rate, seed, offset):
    gen = torch.Generator(device=dy.device)
    gen.manual_seed((seed * 1000003 + offset) & 0x7FFFFFFFFFFFFFFF)
    keep = (torch.rand(dy.shape, generator=gen, device=dy.device) >= rate).to(dy.dtype)
    return dy * keep * (1.0 / (1.0 - rate))


# ----------------------------------------------------------------------------------------------
# losses.  value = mean over batch; gradient pre-scaled by 1/batch (ref loss_ops.cpp)
# ----------------------------------------------------------------------------------------------
def loss_forward_backward(kind, pred, target, param=0.0, want_grad=True):
    """Returns (loss scalar tensor, grad or None, correct-count tensor or None)."""
    b = pred.shape[0]
    p = pred.reshape(b, -1)
    t = target.reshape(b, -1)
    n = p.shape[1]
    correct = None
    nat = _native()
    if nat is not None and p.dtype == torch.float32 and kind in ("softmax_crossentropy", "logsoftmax_crossentropy"):
        loss, grad, correct = nat.softmax_ce(p, t, want_grad)
        return loss, (grad.reshape(pred.shape) if grad is not None else None), correct
    if kind in ("softmax_crossentropy", "logsoftmax_crossentropy"):
        lse = torch.logsumexp(p.double(), dim=1)
        tgt = t.argmax(dim=1)
        loss = (lse - p.double().gather(1, tgt.view(-1, 1)).squeeze(1)).mean().to(p.dtype)
        grad = (torch.softmax(p, dim=1) - t) / b if want_grad else None
        correct = (p.argmax(dim=1) == tgt).sum()
    elif kind == "crossentropy":
        eps = param if param > 0 else 1e-15
        tgt = t.argmax(dim=1)
        pc = p.gather(1, tgt.view(-1, 1)).squeeze(1).clamp(eps, 1 - eps)
        loss = (-torch.log(pc.double())).mean().to(p.dtype)
        grad = (p - t) / b if want_grad else None
        correct = (p.argmax(dim=1) == tgt).sum()
    elif kind == "mse":
        d = p - t
        loss = (d.double() ** 2).mean().to(p.dtype)
        grad = d * (2.0 / (b * n)) if want_grad else None
    elif kind == "mae":
        d = p - t
        loss = d.double().abs().mean().to(p.dtype)
        grad = torch.where(d > 0, 1.0, -1.0).to(p.dtype) / (b * n) if want_grad else None
    elif kind == "huber":
        delta = param
        d = p - t
        a = d.abs()
        l = torch.where(a <= delta, 0.5 * a * a, delta * a - 0.5 * delta * delta)
        loss = l.double().mean().to(p.dtype)
        grad = torch.where(a <= delta, d, delta * torch.sign(d)) / (b * n) if want_grad else None
    else:
        raise ValueError(f"unknown loss {kind}")
    if grad is not None:
        grad = grad.reshape(pred.shape)
    return loss, grad, correct


# ----------------------------------------------------------------------------------------------
# optimizers (multi-tensor semantics: lists of tensors updated in place)
# ----------------------------------------------------------------------------------------------
def _native_ok(tensors):
    return all(t.dtype == torch.float32 and t.is_contiguous() for t in tensors)


def sgd_update(params, grads, velocities, lr, momentum, zero_grads=False, hyper=None):
    nat = _native()
    if nat is not None and _native_ok(list(params) + list(grads)):
        return nat.sgd_update(params, grads, velocities, lr, momentum, zero_grads)
    for i, (p, g) in enumerate(zip(params, grads)):
        if momentum > 0:
            v = velocities[i]
            v.mul_(momentum).add_(g, alpha=-lr)
            p.add_(v)
        else:
            p.add_(g, alpha=-lr)
        if zero_grads:
            g.zero_()


def adam_update(params, grads, ms, vs, lr, beta1, beta2, eps, bc1, bc2, weight_decay, decoupled,
                zero_grads=False, hyper=None):
    """ref adam_kernels.cu:24-60"""
    nat = _native()
    if nat is not None and _native_ok(list(params) + list(grads)):
        return nat.adam_update(params, grads, ms, vs, lr, beta1, beta2, eps, bc1, bc2, weight_decay, decoupled,
                               zero_grads)
    for p, g, m, v in zip(params, grads, ms, vs):
        m.mul_(beta1).add_(g, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        update = lr * (m / bc1) / (torch.sqrt(v / bc2) + eps)
        if weight_decay > 0:
            if decoupled:
                p.mul_(1 - weight_decay * lr)
            else:
                update = update + weight_decay * lr * p
        p.sub_(update)
        if zero_grads:
            g.zero_()


def flatten_nchw(x):
    """[N,C,H,W] -> [N, C*H*W, 1, 1] in NCHW feature order (checkpoint-compatible)."""
    return x.reshape(x.shape[0], -1, 1, 1)


def unflatten_nchw(g, shape):
    return g.reshape(shape)
