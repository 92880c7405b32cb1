"""Functional op set.  Dispatch is by tensor residency only:

* CUDA tensors  -> hand-written sm_100a kernels in ``libdcnn_b200.so`` (``ops.cuda``).  There is no
  silent fallback: if the native library is missing on a GPU box the call raises.
* CPU tensors   -> ATen FP32 (``ops.cpu``), the oracle and the CPU plumbing path.

The op list mirrors the reference's device-dispatched layer ops.
"""
from __future__ import annotations

from . import cpu as _cpu

_cuda = None


def _cuda_backend():
    global _cuda
    if _cuda is None:
        from . import cuda as _c  # raises if the native library cannot be loaded
        _cuda = _c
    return _cuda


def backend_for(t):
    return _cuda_backend() if t.is_cuda else _cpu


def _dispatch(name):
    def f(first, *a, **k):
        t = first[0] if isinstance(first, (list, tuple)) else first
        return getattr(backend_for(t), name)(first, *a, **k)
    f.__name__ = name
    return f


_OPS = [
    "conv2d_forward", "conv2d_backward_data", "conv2d_backward_weight", "bias_grad",
    "dense_forward", "dense_backward_data", "dense_backward_weight",
    "bn_stats", "bn_apply", "bn_backward", "bn_inference",
    "groupnorm_forward", "groupnorm_backward",
    "maxpool_forward", "maxpool_backward", "avgpool_forward", "avgpool_backward",
    "activation_forward", "activation_backward", "add",
    "dropout_forward", "dropout_backward",
    "sgd_update", "adam_update", "flatten_nchw", "unflatten_nchw",
]
for _n in _OPS:
    globals()[_n] = _dispatch(_n)


def bn_finalize(stats, eps, momentum, running_mean, running_var, sync=None, gamma=None, beta=None):
    """-> (mean, invstd).  On CUDA ``invstd`` is an opaque (invstd, scale, shift) bundle: the finalize kernel
    also precomputes the fused apply coefficients from gamma/beta."""
    be = _cpu if isinstance(stats, _cpu.CpuStats) else _cuda_backend()
    return be.bn_finalize(stats, eps, momentum, running_mean, running_var, sync=sync, gamma=gamma, beta=beta)


def loss_forward_backward(kind, pred, target, param=0.0, want_grad=True):
    return backend_for(pred).loss_forward_backward(kind, pred, target, param, want_grad)


def weights_changed():
    """Tell the CUDA backend that parameters were written behind torch's back (a raw-pointer kernel or a CUDA-graph
    replay): cached derived layouts of the weights are rebuilt on their next use.  No-op without the CUDA backend."""
    if _cuda is not None:
        _cuda.weights_changed()


def to_activation_layout(x):
    """Activations on CUDA are physically NHWC (torch channels_last) with a logical NCHW shape."""
    if x.is_cuda and x.dim() == 4:
        return _cuda_backend().as_nhwc(x)
    return x
