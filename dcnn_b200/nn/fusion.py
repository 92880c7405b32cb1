"""Execution units: how containers run their layer lists.

The public contract stays layer-by-layer (every layer still has a working standalone
forward/backward), but ``Sequential`` and ``ResidualBlock`` execute *units*.  A unit is either
one layer or a fused group:

  ConvBNActUnit = Conv2D -> BatchNorm -> [ReLU]
  GemmReluUnit  = Conv2D | Dense -> ReLU              (ReLU in the GEMM epilogue)

which runs as  conv (BN partial statistics produced by the conv epilogue)  ->  finalize  ->
one apply pass (scale/shift [+residual] [+ReLU]).  The reference runs the same chain as
cudnnConvolutionForward + cudnnAddTensor + fused_stats + fused_apply + copy + copy + relu
(SURVEY §3.1), i.e. ~3x the HBM traffic.  Backward is symmetric: one reduce pass + one apply
pass (ReLU mask folded in), then dgrad/wgrad.
"""
from __future__ import annotations

from typing import List, Optional

from .. import ops
from .layers import ActivationLayer, BatchNormLayer, Conv2DLayer, DenseLayer, Layer


class SingleUnit:
    def __init__(self, layer: Layer):
        self.layers = [layer]
        self.layer = layer

    @property
    def name(self):
        return self.layer.name

    def forward(self, x, mb, residual=None, relu=False):
        assert residual is None and not relu
        return self.layer.forward(x, mb)

    def backward(self, g, mb, want_masked=False, accumulate_into=None):
        if isinstance(self.layer, Conv2DLayer) and accumulate_into is not None:
            return self.layer.backward(g, mb, out=accumulate_into, accumulate=True), None
        dx = self.layer.backward(g, mb)
        if accumulate_into is not None and dx is not None:
            dx = ops.add(dx, accumulate_into)
        return dx, None

    def can_fuse_residual(self):
        return False


class ConvBNActUnit:
    def __init__(self, conv: Optional[Conv2DLayer], bn: BatchNormLayer, act: Optional[ActivationLayer]):
        self.conv, self.bn, self.act = conv, bn, act
        self.layers = [l for l in (conv, bn, act) if l is not None]

    @property
    def name(self):
        return "+".join(l.name for l in self.layers)

    def can_fuse_residual(self):
        return self.act is None

    def forward(self, x, mb, residual=None, relu=False):
        relu = relu or self.act is not None
        stats = None
        if self.conv is not None:
            x, stats = self.conv.forward_with_stats(x, mb, want_stats=self.bn.training)
        return self.bn.forward_fused(x, stats, mb, relu=relu, residual=residual)

    def backward(self, g, mb, want_masked=False, accumulate_into=None):
        dy, masked = self.bn.backward_fused(g, mb, want_masked=want_masked)
        if self.conv is None:
            if accumulate_into is not None:
                dy = ops.add(dy, accumulate_into)
            return dy, masked
        if accumulate_into is not None:
            dx = self.conv.backward(dy, mb, out=accumulate_into, accumulate=True)
        else:
            dx = self.conv.backward(dy, mb)
        return dx, masked


class GemmReluUnit:
    """Conv2D / Dense -> ReLU without a BatchNorm in between: the ReLU runs in the GEMM epilogue (SURVEY K10 / K22:
    "epilogue bias(+ReLU)"), so the pre-activation tensor is never written.  The activation layer caches the OUTPUT
    instead of its input -- for ReLU the mask (out > 0) is the same -- and backward is the two layers' own."""

    def __init__(self, gemm: Layer, act: ActivationLayer):
        self.gemm, self.act = gemm, act
        self.layers = [gemm, act]

    @property
    def name(self):
        return self.gemm.name + "+" + self.act.name

    def can_fuse_residual(self):
        return False

    def forward(self, x, mb, residual=None, relu=False):
        assert residual is None and not relu
        y = self.gemm.forward(x, mb, relu=True)
        if self.act.training:
            self.act._cache[mb] = y
        return y

    def backward(self, g, mb, want_masked=False, accumulate_into=None):
        g = self.act.backward(g, mb)
        if isinstance(self.gemm, Conv2DLayer) and accumulate_into is not None:
            return self.gemm.backward(g, mb, out=accumulate_into, accumulate=True), None
        dx = self.gemm.backward(g, mb)
        if accumulate_into is not None and dx is not None:
            dx = ops.add(dx, accumulate_into)
        return dx, None


def _is_relu(layer):
    return isinstance(layer, ActivationLayer) and layer.activation.name() == "relu"


def build_units(layers: List[Layer], fuse: bool = True):
    """Greedy peephole grouping of a layer list into execution units."""
    units = []
    i, n = 0, len(layers)
    while i < n:
        l = layers[i]
        if fuse and isinstance(l, Conv2DLayer) and i + 1 < n and isinstance(layers[i + 1], BatchNormLayer) \
                and layers[i + 1].num_features == l.out_channels:
            bn = layers[i + 1]
            act = layers[i + 2] if i + 2 < n and _is_relu(layers[i + 2]) else None
            units.append(ConvBNActUnit(l, bn, act))
            i += 3 if act is not None else 2
        elif fuse and isinstance(l, (Conv2DLayer, DenseLayer)) and i + 1 < n and _is_relu(layers[i + 1]):
            units.append(GemmReluUnit(l, layers[i + 1]))
            i += 2
        elif fuse and isinstance(l, BatchNormLayer):
            act = layers[i + 1] if i + 1 < n and _is_relu(layers[i + 1]) else None
            units.append(ConvBNActUnit(None, l, act))
            i += 2 if act is not None else 1
        else:
            units.append(SingleUnit(l))
            i += 1
    return units
