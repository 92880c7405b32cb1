"""Layers with the reference's explicit per-microbatch forward/backward contract.

Mirrors ``Layer<T>`` (/root/reference/include/nn/layers_impl/base_layer.hpp:37-116) and the nine
concrete layers (include/nn/layers_impl/*.hpp): every layer caches what its backward needs in a
dict keyed by ``micro_batch_id`` -- that is the whole microbatching mechanism.  There is no
autograd anywhere in this package.

B200-specific design: on CUDA, activations are logical NCHW tensors stored channels-last (NHWC),
conv weights are logical OIHW stored O-H-W-I, so the hand-written implicit-GEMM kernels see
K-major operands without any transposition.  In addition to the standalone ``forward/backward``
each layer exposes ``*_fused`` entry points that the containers (``Sequential``,
``ResidualBlock``) use to run conv -> BN(+stats in the conv epilogue) -> ReLU(+residual add) as
two kernels instead of six.
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

from .. import ops

UINT32_MAX = 0xFFFFFFFF


def _numel(shape):
    n = 1
    for s in shape:
        n *= int(s)
    return n


def _pair(v):
    return (int(v[0]), int(v[1])) if isinstance(v, (tuple, list)) else (int(v), int(v))


class Layer:
    """Base class. Subclasses implement forward/backward/compute_output_shape/get_config."""

    type_name = "layer"

    def __init__(self, name: str = ""):
        self.name = name
        self.training = True
        self.device = torch.device("cpu")
        self.seed: Optional[int] = None
        self.enable_profiling_ = False
        self.perf_timers: Dict[str, float] = {}
        self._initialized = False

    # -- lifecycle ---------------------------------------------------------------------------
    def initialize(self):
        self._initialized = True

    def set_seed(self, seed: int):
        self.seed = int(seed)

    def set_training(self, training: bool):
        self.training = bool(training)

    def is_training(self) -> bool:
        return self.training

    def set_device(self, device):
        self.device = torch.device(device)
        self._move_to_device()

    def get_device(self):
        return self.device

    def _move_to_device(self):
        pass

    # -- compute -----------------------------------------------------------------------------
    def forward(self, x: torch.Tensor, micro_batch_id: int = 0) -> torch.Tensor:
        raise NotImplementedError

    def backward(self, grad: torch.Tensor, micro_batch_id: int = 0) -> torch.Tensor:
        raise NotImplementedError

    def __call__(self, x, micro_batch_id: int = 0):
        return self.forward(x, micro_batch_id)

    # -- parameters --------------------------------------------------------------------------
    def parameters(self) -> List[torch.Tensor]:
        return []

    def gradients(self) -> List[torch.Tensor]:
        return []

    def buffers(self) -> Dict[str, torch.Tensor]:
        """Non-parameter state (BN running stats). Not part of the reference .bin; goes to the
        optional ``.state`` sidecar."""
        return {}

    def has_parameters(self) -> bool:
        return False

    #: attribute names behind parameters()/gradients(), in order (used to re-home them into flat buffers)
    _param_attrs: tuple = ()
    _grad_attrs: tuple = ()

    def rebind_parameters(self, params, grads):
        """Replace this layer's parameter/gradient tensors by same-shaped views (e.g. slices of a flat,
        NVLink-shared buffer).  Values are copied into the new storage.  Returns the number consumed."""
        live = [(a, g) for a, g in zip(self._param_attrs, self._grad_attrs) if getattr(self, a) is not None]
        for i, (a, g) in enumerate(live):
            params[i].copy_(getattr(self, a))
            grads[i].copy_(getattr(self, g))
            setattr(self, a, params[i])
            setattr(self, g, grads[i])
        return len(live)

    def clear_gradients(self):
        for g in self.gradients():
            g.zero_()

    # -- metadata ----------------------------------------------------------------------------
    def type(self) -> str:
        return self.type_name

    def get_config(self) -> Dict[str, Any]:
        return {"name": self.name, "parameters": {}}

    def clone(self) -> "Layer":
        from .factory import LayerFactory
        cfg = self.get_config()
        new = LayerFactory.create(self.type(), cfg)
        return new

    def compute_output_shape(self, input_shape: Sequence[int]) -> List[int]:
        return list(input_shape)

    def forward_flops(self, input_shape) -> int:
        return 0

    def backward_flops(self, input_shape) -> int:
        return 0

    def forward_complexity(self, input_shape) -> int:
        return min(self.forward_flops(input_shape), UINT32_MAX)

    def backward_complexity(self, input_shape) -> int:
        return min(self.backward_flops(input_shape), UINT32_MAX)

    def cached_memory_bytes(self) -> int:
        return 0

    def clear_cache(self):
        pass

    def enable_profiling(self, enable: bool = True):
        self.enable_profiling_ = enable

    def reset_profiling_info(self):
        self.perf_timers.clear()

    def print_profiling_info(self):
        print(f"Profiling info for layer: {self.name}")
        for k, v in self.perf_timers.items():
            print(f"  {k}: {v} ms")

    # helpers
    def _rng(self):
        g = torch.Generator(device="cpu")
        if self.seed is not None:
            g.manual_seed(self.seed)
        else:
            g.seed()
        return g

    def _to_dev(self, t: torch.Tensor, weight_4d: bool = False) -> torch.Tensor:
        t = t.to(self.device)
        if weight_4d and t.is_cuda and t.dim() == 4:
            t = t.contiguous(memory_format=torch.channels_last)
        return t


def _cache_bytes(*dicts):
    total = 0
    for d in dicts:
        for v in d.values():
            items = v if isinstance(v, (tuple, list)) else (v,)
            for t in items:
                if isinstance(t, torch.Tensor):
                    total += t.numel() * t.element_size()
    return total


# ==============================================================================================
class Conv2DLayer(Layer):
    """2-D convolution (ref include/nn/layers_impl/conv2d_layer.{hpp,tpp}).

    Weights ``[OC, IC, KH, KW]``, bias ``[OC,1,1,1]``; init ``U(-1/sqrt(fan_in), 1/sqrt(fan_in))``
    (conv2d_layer.tpp:70-86).  Weight/bias gradients accumulate across microbatches (beta=1,
    cudnn_conv2d_ops.cu:226,242)."""

    type_name = "conv2d"
    _param_attrs = ("weights", "bias")
    _grad_attrs = ("weight_grad", "bias_grad")

    def __init__(self, in_channels, out_channels, kernel_h, kernel_w, stride_h=1, stride_w=1,
                 pad_h=0, pad_w=0, use_bias=True, name="conv2d"):
        super().__init__(name)
        self.in_channels, self.out_channels = int(in_channels), int(out_channels)
        self.kernel = (int(kernel_h), int(kernel_w))
        self.stride = (int(stride_h), int(stride_w))
        self.pad = (int(pad_h), int(pad_w))
        self.use_bias = bool(use_bias)
        self.weights = self.bias = self.weight_grad = self.bias_grad = None
        self._inputs: Dict[int, torch.Tensor] = {}
        self.need_input_grad = True

    def initialize(self):
        if self._initialized:
            return
        fan_in = self.in_channels * self.kernel[0] * self.kernel[1]
        bound = 1.0 / math.sqrt(fan_in)
        g = self._rng()
        w = (torch.rand(self.out_channels, self.in_channels, *self.kernel, generator=g) * 2 - 1) * bound
        self.weights = self._to_dev(w, True)
        self.weight_grad = torch.zeros_like(self.weights)
        if self.use_bias:
            b = (torch.rand(self.out_channels, 1, 1, 1, generator=g) * 2 - 1) * bound
            self.bias = self._to_dev(b)
            self.bias_grad = torch.zeros_like(self.bias)
        self._initialized = True

    def _move_to_device(self):
        if self.weights is not None:
            self.weights = self._to_dev(self.weights, True)
            self.weight_grad = self._to_dev(self.weight_grad, True)
            if self.use_bias:
                self.bias = self._to_dev(self.bias)
                self.bias_grad = self._to_dev(self.bias_grad)

    def forward(self, x, micro_batch_id=0, relu=False):
        y, _ = self.forward_with_stats(x, micro_batch_id, want_stats=False, relu=relu)
        return y

    def forward_with_stats(self, x, micro_batch_id=0, want_stats=True, relu=False):
        if not self._initialized:
            self.initialize()
        if x.device != self.weights.device:
            x = x.to(self.weights.device)
        x = ops.to_activation_layout(x)
        if x.shape[1] != self.in_channels:
            raise ValueError(f"{self.name}: expected {self.in_channels} input channels, got {x.shape[1]}")
        if self.training:
            self._inputs[micro_batch_id] = x
        return ops.conv2d_forward(x, self.weights, self.bias, self.stride, self.pad, want_stats, relu)

    def backward(self, grad, micro_batch_id=0, out=None, accumulate=False):
        x = self._inputs.get(micro_batch_id)
        if x is None:
            raise RuntimeError(f"{self.name}: no cached input for micro-batch {micro_batch_id}")
        grad = ops.to_activation_layout(grad)
        # the bias gradient rides in the weight-gradient launch (K13: the reference calls cudnnConvolutionBackwardBias)
        ops.conv2d_backward_weight(grad, x, self.weight_grad, self.stride, self.pad,
                                   self.bias_grad if self.use_bias else None)
        if not self.need_input_grad:
            return None
        return ops.conv2d_backward_data(grad, self.weights, x.shape, self.stride, self.pad,
                                        out=out, accumulate=accumulate)

    def parameters(self):
        return [self.weights, self.bias] if self.use_bias else [self.weights]

    def gradients(self):
        return [self.weight_grad, self.bias_grad] if self.use_bias else [self.weight_grad]

    def has_parameters(self):
        return True

    def get_config(self):
        return {"name": self.name, "parameters": {
            "in_channels": self.in_channels, "out_channels": self.out_channels,
            "kernel_h": self.kernel[0], "kernel_w": self.kernel[1],
            "stride_h": self.stride[0], "stride_w": self.stride[1],
            "pad_h": self.pad[0], "pad_w": self.pad[1],
            "use_bias": self.use_bias, "optimized": "tcgen05"}}

    def compute_output_shape(self, s):
        if len(s) != 4:
            raise ValueError("Conv2D requires 4D input")
        oh = (s[2] + 2 * self.pad[0] - self.kernel[0]) // self.stride[0] + 1
        ow = (s[3] + 2 * self.pad[1] - self.kernel[1]) // self.stride[1] + 1
        return [s[0], self.out_channels, oh, ow]

    def forward_flops(self, s):
        o = self.compute_output_shape(s)
        out_size = o[0] * o[2] * o[3]
        ksz = self.in_channels * self.kernel[0] * self.kernel[1]
        return 2 * self.out_channels * ksz * out_size + (out_size * self.out_channels if self.use_bias else 0)

    def backward_flops(self, s):
        o = self.compute_output_shape(s)
        out_size = o[0] * o[2] * o[3]
        ksz = self.in_channels * self.kernel[0] * self.kernel[1]
        return 4 * self.out_channels * ksz * out_size + (out_size * self.out_channels if self.use_bias else 0)

    def cached_memory_bytes(self):
        return _cache_bytes(self._inputs)

    def clear_cache(self):
        self._inputs.clear()


# ==============================================================================================
class DenseLayer(Layer):
    """Fully connected layer (ref dense_layer.{hpp,tpp}). Accepts any ``[N,C,H,W]`` with
    ``C*H*W == input_features`` and emits ``[N,out,1,1]`` (dense_layer.tpp:64-88)."""

    type_name = "dense"
    _param_attrs = ("weights", "bias")
    _grad_attrs = ("weight_grad", "bias_grad")

    def __init__(self, input_features, output_features, use_bias=True, name="dense"):
        super().__init__(name)
        self.input_features, self.output_features = int(input_features), int(output_features)
        self.use_bias = bool(use_bias)
        self.weights = self.bias = self.weight_grad = self.bias_grad = None
        self._inputs: Dict[int, Tuple[torch.Tensor, Tuple[int, ...]]] = {}
        self.need_input_grad = True

    def initialize(self):
        if self._initialized:
            return
        bound = 1.0 / math.sqrt(self.input_features)
        g = self._rng()
        w = (torch.rand(self.output_features, self.input_features, 1, 1, generator=g) * 2 - 1) * bound
        self.weights = self._to_dev(w).contiguous()
        self.weight_grad = torch.zeros_like(self.weights)
        if self.use_bias:
            b = (torch.rand(self.output_features, 1, 1, 1, generator=g) * 2 - 1) * bound
            self.bias = self._to_dev(b)
            self.bias_grad = torch.zeros_like(self.bias)
        self._initialized = True

    def _move_to_device(self):
        if self.weights is not None:
            self.weights = self.weights.to(self.device).contiguous()
            self.weight_grad = self.weight_grad.to(self.device).contiguous()
            if self.use_bias:
                self.bias = self.bias.to(self.device)
                self.bias_grad = self.bias_grad.to(self.device)

    def forward(self, x, micro_batch_id=0, relu=False):
        if not self._initialized:
            self.initialize()
        if x.device != self.weights.device:
            x = x.to(self.weights.device)
        shape = tuple(x.shape)
        if _numel(shape[1:]) != self.input_features:
            raise ValueError(f"{self.name}: expected {self.input_features} features, got {shape}")
        x2 = ops.flatten_nchw(x).reshape(shape[0], self.input_features)
        if self.training:
            self._inputs[micro_batch_id] = (x2, shape)
        y = ops.dense_forward(x2, self.weights.view(self.output_features, self.input_features),
                              self.bias, relu)
        return y.reshape(shape[0], self.output_features, 1, 1)

    def backward(self, grad, micro_batch_id=0):
        x2, shape = self._inputs[micro_batch_id]
        g2 = grad.reshape(shape[0], self.output_features)
        w2 = self.weights.view(self.output_features, self.input_features)
        ops.dense_backward_weight(g2, x2, self.weight_grad.view_as(w2), self.bias_grad if self.use_bias else None)
        if not self.need_input_grad:
            return None
        dx = ops.dense_backward_data(g2, w2)
        return ops.unflatten_nchw(dx.reshape(shape[0], self.input_features, 1, 1), shape)

    def parameters(self):
        return [self.weights, self.bias] if self.use_bias else [self.weights]

    def gradients(self):
        return [self.weight_grad, self.bias_grad] if self.use_bias else [self.weight_grad]

    def has_parameters(self):
        return True

    def get_config(self):
        return {"name": self.name, "parameters": {
            "input_features": self.input_features, "output_features": self.output_features,
            "use_bias": self.use_bias, "optimized": "native"}}

    def compute_output_shape(self, s):
        return [s[0], self.output_features, 1, 1]

    def forward_flops(self, s):
        return 2 * s[0] * self.input_features * self.output_features + (s[0] * self.output_features if self.use_bias else 0)

    def backward_flops(self, s):
        return 4 * s[0] * self.input_features * self.output_features + (s[0] * self.output_features if self.use_bias else 0)

    def cached_memory_bytes(self):
        return sum(v[0].numel() * 4 for v in self._inputs.values())

    def clear_cache(self):
        self._inputs.clear()


# ==============================================================================================
class BatchNormLayer(Layer):
    """Batch normalisation over (N,H,W) per channel (ref batchnorm_layer.{hpp,tpp},
    batchnorm_ops.cu:108-295).  Statistics are per *microbatch*; biased variance normalises,
    unbiased variance feeds the running estimate; ``running = (1-m)*running + m*batch``.

    ``sync`` (set by the data-parallel wrapper) makes the statistics cross-replica: the finalize
    kernel then reads every replica's partial sums over NVLink peer memory (K27)."""

    type_name = "batchnorm"
    _param_attrs = ("gamma", "beta")
    _grad_attrs = ("gamma_grad", "beta_grad")

    def __init__(self, num_features, epsilon=1e-5, momentum=0.1, affine=True, name="batchnorm"):
        super().__init__(name)
        self.num_features = int(num_features)
        self.epsilon, self.momentum, self.affine = float(epsilon), float(momentum), bool(affine)
        self.gamma = self.beta = self.gamma_grad = self.beta_grad = None
        self.running_mean = self.running_var = None
        self._cache: Dict[int, tuple] = {}
        self.sync = None

    def initialize(self):
        if self._initialized:
            return
        c = self.num_features
        if self.affine:
            self.gamma = torch.ones(c, 1, 1, 1, device=self.device)
            self.beta = torch.zeros(c, 1, 1, 1, device=self.device)
            self.gamma_grad = torch.zeros_like(self.gamma)
            self.beta_grad = torch.zeros_like(self.beta)
        self.running_mean = torch.zeros(c, device=self.device)
        self.running_var = torch.ones(c, device=self.device)
        self._initialized = True

    def _move_to_device(self):
        for n in ("gamma", "beta", "gamma_grad", "beta_grad", "running_mean", "running_var"):
            t = getattr(self, n)
            if t is not None:
                setattr(self, n, t.to(self.device))

    def forward(self, x, micro_batch_id=0):
        return self.forward_fused(x, None, micro_batch_id, relu=False, residual=None)

    def forward_fused(self, x, stats, micro_batch_id=0, relu=False, residual=None):
        """BN (+residual add) (+ReLU) in one apply pass.  ``stats`` are the per-channel partial
        statistics already produced by the upstream conv's epilogue (or None)."""
        if not self._initialized:
            self.initialize()
        if x.device != self.running_mean.device:
            x = x.to(self.running_mean.device)
        x = ops.to_activation_layout(x)
        if x.shape[1] != self.num_features:
            raise ValueError(f"{self.name}: expected {self.num_features} channels, got {x.shape[1]}")
        if not self.training:
            return ops.bn_inference(x, self.running_mean, self.running_var, self.gamma, self.beta,
                                    self.epsilon, residual, relu)
        if stats is None:
            stats = ops.bn_stats(x)
        mean, invstd = ops.bn_finalize(stats, self.epsilon, self.momentum, self.running_mean,
                                       self.running_var, sync=self.sync, gamma=self.gamma, beta=self.beta)
        mask = None
        if relu and residual is not None and x.is_cuda and x.shape[1] % 4 == 0 and ops._cuda_backend().RELU_BITMASK:
            # compact ReLU mask (1/16 of the tensor) so backward does not re-read `out` twice
            mask = torch.empty(x.numel() // 4, dtype=torch.uint8, device=x.device)
            out = ops.bn_apply(x, mean, invstd, self.gamma, self.beta, residual, relu, mask_out=mask)
        else:
            out = ops.bn_apply(x, mean, invstd, self.gamma, self.beta, residual, relu)
        keep_out = out if (relu and mask is None) else None
        self._cache[micro_batch_id] = (x, mean, invstd, keep_out, relu, relu and residual is None, mask)
        return out

    def backward(self, grad, micro_batch_id=0):
        dx, _ = self.backward_fused(grad, micro_batch_id, want_masked=False)
        return dx

    def backward_fused(self, grad, micro_batch_id=0, want_masked=False):
        x, mean, invstd, out, relu, plain_relu, mask = self._cache[micro_batch_id]
        grad = ops.to_activation_layout(grad)
        # plain_relu: no residual between BN and ReLU, so the mask can be recomputed from x (CUDA backend)
        return ops.bn_backward(grad, x, out, mean, invstd, self.gamma, self.gamma_grad,
                               self.beta_grad, relu=relu, want_masked=want_masked, sync=self.sync,
                               recompute_mask=plain_relu, relu_mask=mask)

    def parameters(self):
        return [self.gamma, self.beta] if self.affine else []

    def gradients(self):
        return [self.gamma_grad, self.beta_grad] if self.affine else []

    def buffers(self):
        return {"running_mean": self.running_mean, "running_var": self.running_var}

    def has_parameters(self):
        return self.affine

    def get_config(self):
        return {"name": self.name, "parameters": {
            "num_features": self.num_features, "epsilon": self.epsilon,
            "momentum": self.momentum, "affine": self.affine}}

    def forward_flops(self, s):
        n = _numel(s)
        mean_f = n
        return mean_f + (2 * n + mean_f) + 3 * n + (2 * n if self.affine else 0)

    def backward_flops(self, s):
        n = _numel(s)
        return (2 * n if self.affine else 0) + 9 * n

    def cached_memory_bytes(self):
        return _cache_bytes(self._cache)

    def clear_cache(self):
        self._cache.clear()


# ==============================================================================================
class GroupNormLayer(Layer):
    """Group normalisation (ref groupnorm_layer.{hpp,tpp}, groupnorm_ops.cu:45-252)."""

    type_name = "groupnorm"
    _param_attrs = ("gamma", "beta")
    _grad_attrs = ("gamma_grad", "beta_grad")

    def __init__(self, num_groups, num_channels, epsilon=1e-5, affine=True, name="groupnorm"):
        super().__init__(name)
        self.num_groups, self.num_channels = int(num_groups), int(num_channels)
        if self.num_channels % self.num_groups:
            raise ValueError("num_channels must be divisible by num_groups")
        self.epsilon, self.affine = float(epsilon), bool(affine)
        self.gamma = self.beta = self.gamma_grad = self.beta_grad = None
        self._cache: Dict[int, tuple] = {}

    def initialize(self):
        if self._initialized:
            return
        if self.affine:
            self.gamma = torch.ones(self.num_channels, 1, 1, 1, device=self.device)
            self.beta = torch.zeros(self.num_channels, 1, 1, 1, device=self.device)
            self.gamma_grad = torch.zeros_like(self.gamma)
            self.beta_grad = torch.zeros_like(self.beta)
        self._initialized = True

    def _move_to_device(self):
        for n in ("gamma", "beta", "gamma_grad", "beta_grad"):
            t = getattr(self, n)
            if t is not None:
                setattr(self, n, t.to(self.device))

    def forward(self, x, micro_batch_id=0):
        if not self._initialized:
            self.initialize()
        x = ops.to_activation_layout(x)
        y, mean, invstd = ops.groupnorm_forward(x, self.num_groups, self.gamma, self.beta, self.epsilon)
        if self.training:
            self._cache[micro_batch_id] = (x, mean, invstd)
        return y

    def backward(self, grad, micro_batch_id=0):
        x, mean, invstd = self._cache[micro_batch_id]
        grad = ops.to_activation_layout(grad)
        return ops.groupnorm_backward(grad, x, mean, invstd, self.num_groups, self.gamma,
                                      self.gamma_grad, self.beta_grad)

    def parameters(self):
        return [self.gamma, self.beta] if self.affine else []

    def gradients(self):
        return [self.gamma_grad, self.beta_grad] if self.affine else []

    def has_parameters(self):
        return self.affine

    def get_config(self):
        return {"name": self.name, "parameters": {
            "num_groups": self.num_groups, "num_channels": self.num_channels,
            "epsilon": self.epsilon, "affine": self.affine}}

    def forward_flops(self, s):
        n = _numel(s)
        return n + (2 * n + n) + 3 * n + (2 * n if self.affine else 0)

    def backward_flops(self, s):
        n = _numel(s)
        return (2 * n if self.affine else 0) + 9 * n

    def cached_memory_bytes(self):
        return _cache_bytes(self._cache)

    def clear_cache(self):
        self._cache.clear()


# ==============================================================================================
class _Pool2D(Layer):
    def __init__(self, pool_h, pool_w, stride_h=0, stride_w=0, pad_h=0, pad_w=0, name="pool"):
        super().__init__(name)
        self.pool = (int(pool_h), int(pool_w))
        # stride 0 => pool size (ref maxpool2d_layer.tpp:21)
        self.stride = (int(stride_h) or int(pool_h), int(stride_w) or int(pool_w))
        self.pad = (int(pad_h), int(pad_w))
        self._cache: Dict[int, tuple] = {}

    def get_config(self):
        return {"name": self.name, "parameters": {
            "pool_h": self.pool[0], "pool_w": self.pool[1],
            "stride_h": self.stride[0], "stride_w": self.stride[1],
            "pad_h": self.pad[0], "pad_w": self.pad[1]}}

    def compute_output_shape(self, s):
        oh = (s[2] + 2 * self.pad[0] - self.pool[0]) // self.stride[0] + 1
        ow = (s[3] + 2 * self.pad[1] - self.pool[1]) // self.stride[1] + 1
        return [s[0], s[1], oh, ow]

    def cached_memory_bytes(self):
        return _cache_bytes(self._cache)

    def clear_cache(self):
        self._cache.clear()


class MaxPool2DLayer(_Pool2D):
    """Max pooling with implicit -inf padding; ties resolve to the first max in scan order
    (ref maxpool_ops.cu:15-82).  The CUDA path stores a 1-byte window index, not an 8-byte mask."""

    type_name = "maxpool2d"

    def __init__(self, pool_h, pool_w, stride_h=0, stride_w=0, pad_h=0, pad_w=0, name="maxpool2d"):
        super().__init__(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w, name)

    def forward(self, x, micro_batch_id=0):
        x = ops.to_activation_layout(x)
        y, idx = ops.maxpool_forward(x, self.pool, self.stride, self.pad)
        if self.training:
            self._cache[micro_batch_id] = (idx, tuple(x.shape))
        return y

    def backward(self, grad, micro_batch_id=0):
        idx, shape = self._cache[micro_batch_id]
        grad = ops.to_activation_layout(grad)
        return ops.maxpool_backward(grad, idx, shape, self.pool, self.stride, self.pad)

    def forward_flops(self, s):
        o = self.compute_output_shape(s)
        return _numel(o) * self.pool[0] * self.pool[1]

    def backward_flops(self, s):
        return _numel(self.compute_output_shape(s))


class AvgPool2DLayer(_Pool2D):
    """Average pooling; divisor is always pool_h*pool_w (ref avgpool_ops.cu:106,122)."""

    type_name = "avgpool2d"

    def __init__(self, pool_h, pool_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0, name="avgpool2d"):
        super().__init__(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w, name)

    def forward(self, x, micro_batch_id=0):
        x = ops.to_activation_layout(x)
        if self.training:
            self._cache[micro_batch_id] = (tuple(x.shape),)
        return ops.avgpool_forward(x, self.pool, self.stride, self.pad)

    def backward(self, grad, micro_batch_id=0):
        (shape,) = self._cache[micro_batch_id]
        grad = ops.to_activation_layout(grad)
        return ops.avgpool_backward(grad, shape, self.pool, self.stride, self.pad)

    def forward_flops(self, s):
        o = self.compute_output_shape(s)
        return _numel(o) * (self.pool[0] * self.pool[1] + 1)

    def backward_flops(self, s):
        return 2 * _numel(s)


# ==============================================================================================
class DropoutLayer(Layer):
    """Inverted dropout (ref dropout_layer.{hpp,tpp}, dropout.cu:28-125).  Counter-based Philox:
    the mask is regenerated in backward from (seed, per-call offset), no mask tensor is stored."""

    type_name = "dropout"

    def __init__(self, dropout_rate, name="dropout"):
        super().__init__(name)
        self.dropout_rate = float(dropout_rate)
        if not 0.0 <= self.dropout_rate < 1.0:
            raise ValueError("dropout_rate must be in [0,1)")
        self._calls = 0
        self._cache: Dict[int, int] = {}

    def forward(self, x, micro_batch_id=0):
        if not self.training or self.dropout_rate == 0.0:
            return x
        seed = self.seed if self.seed is not None else 0x5EED
        self._calls += 1
        self._cache[micro_batch_id] = self._calls
        return ops.dropout_forward(x, self.dropout_rate, seed, self._calls)

    def backward(self, grad, micro_batch_id=0):
        if not self.training or self.dropout_rate == 0.0:
            return grad
        seed = self.seed if self.seed is not None else 0x5EED
        return ops.dropout_backward(grad, self.dropout_rate, seed, self._cache[micro_batch_id])

    def get_config(self):
        return {"name": self.name, "parameters": {"dropout_rate": self.dropout_rate}}

    def forward_flops(self, s):
        if not self.training:
            return 0
        n = _numel(s)
        return 2 * n + int((1.0 - self.dropout_rate) * n)

    def backward_flops(self, s):
        return _numel(s) if self.training else 0


class FlattenLayer(Layer):
    """[N,C,H,W] -> [N, C*H*W, 1, 1] in NCHW feature order (ref flatten_layer.tpp)."""

    type_name = "flatten"

    def __init__(self, name="flatten"):
        super().__init__(name)
        self._cache: Dict[int, Tuple[int, ...]] = {}

    def forward(self, x, micro_batch_id=0):
        self._cache[micro_batch_id] = tuple(x.shape)
        return ops.flatten_nchw(x)

    def backward(self, grad, micro_batch_id=0):
        return ops.unflatten_nchw(grad, self._cache[micro_batch_id])

    def compute_output_shape(self, s):
        return [s[0], _numel(s[1:]), 1, 1]

    def clear_cache(self):
        self._cache.clear()


class ActivationLayer(Layer):
    """Elementwise / channel-softmax activation as a layer (ref activation_layer.tpp:22-60).
    Gradients are recomputed from the cached pre-activation input."""

    type_name = "activation"

    def __init__(self, activation, name="activation"):
        super().__init__(name)
        from .activations import create_activation, ActivationFunction
        self.activation = activation if isinstance(activation, ActivationFunction) else create_activation(activation)
        if self.activation is None:
            raise ValueError("Failed to create activation")
        self._cache: Dict[int, torch.Tensor] = {}

    def forward(self, x, micro_batch_id=0):
        if self.training:
            self._cache[micro_batch_id] = x
        return self.activation.forward(x)

    def backward(self, grad, micro_batch_id=0):
        return self.activation.backward(self._cache[micro_batch_id], grad)

    def get_config(self):
        return {"name": self.name, "parameters": {"activation": self.activation.name()}}

    def forward_flops(self, s):
        return 2 * _numel(s)

    def backward_flops(self, s):
        return 2 * _numel(s)

    def cached_memory_bytes(self):
        return _cache_bytes(self._cache)

    def clear_cache(self):
        self._cache.clear()
