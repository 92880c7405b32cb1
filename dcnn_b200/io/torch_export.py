"""Sequential <-> plain ``torch.nn.Module``.

``to_torch_module(model)`` builds the equivalent module tree out of stock PyTorch layers (Conv2d, BatchNorm2d,
GroupNorm, MaxPool2d, AvgPool2d, Linear, Dropout, the activations, a residual wrapper) and copies parameters and
BatchNorm statistics into it; ``load_from_torch_module`` copies them back.  Uses: hand a trained network to code that
expects an ``nn.Module`` (export, autograd-based analysis, TorchScript / ONNX), and run the same architecture through
PyTorch's own kernels (cuDNN / cuBLAS) as a second baseline -- the role of the reference's ``torch/`` scripts
(torch/torch_resnet9_cifar10.py, torch_tiny_imagenet_trainer.py, resnet_modules.py), see
``benchmarks/torch_baseline.py``.  Outputs keep the reference's 4-D convention: Flatten yields [N, F, 1, 1] and Dense
maps [N, F, 1, 1] -> [N, O, 1, 1]."""
from __future__ import annotations

from typing import List

import torch
from torch import nn as tnn

from ..nn.blocks import ResidualBlock
from ..nn.layers import (ActivationLayer, AvgPool2DLayer, BatchNormLayer, Conv2DLayer, DenseLayer, DropoutLayer,
                         FlattenLayer, GroupNormLayer, MaxPool2DLayer)


class Flatten4D(tnn.Module):
    def forward(self, x):
        return x.reshape(x.shape[0], -1, 1, 1)


class Dense4D(tnn.Linear):
    """Linear over [N, F, 1, 1] (or any [N, ...] with F elements per sample), output [N, O, 1, 1]."""

    def forward(self, x):
        return super().forward(x.reshape(x.shape[0], -1)).reshape(x.shape[0], -1, 1, 1)


class ChannelSoftmax(tnn.Module):
    def forward(self, x):
        return torch.softmax(x, dim=1)


class Residual(tnn.Module):
    def __init__(self, main: tnn.Module, shortcut: tnn.Module, activation: tnn.Module):
        super().__init__()
        self.main, self.shortcut, self.activation = main, shortcut, activation

    def forward(self, x):
        return self.activation(self.main(x) + self.shortcut(x))


def _activation(fn) -> tnn.Module:
    kind = fn.kind if hasattr(fn, "kind") else str(fn)
    if kind == "relu":
        return tnn.ReLU()
    if kind == "leaky_relu":
        return tnn.LeakyReLU(fn.alpha)
    if kind == "elu":
        return tnn.ELU(fn.alpha)
    if kind == "sigmoid":
        return tnn.Sigmoid()
    if kind == "tanh":
        return tnn.Tanh()
    if kind == "softmax":
        return ChannelSoftmax()
    if kind in ("linear", "none", ""):
        return tnn.Identity()
    raise ValueError(f"no torch counterpart for activation '{kind}'")


def _convert(layer) -> tnn.Module:
    if isinstance(layer, Conv2DLayer):
        return tnn.Conv2d(layer.in_channels, layer.out_channels, layer.kernel, layer.stride, layer.pad,
                          bias=layer.use_bias)
    if isinstance(layer, DenseLayer):
        return Dense4D(layer.input_features, layer.output_features, bias=layer.use_bias)
    if isinstance(layer, BatchNormLayer):
        return tnn.BatchNorm2d(layer.num_features, eps=layer.epsilon, momentum=layer.momentum, affine=layer.affine)
    if isinstance(layer, GroupNormLayer):
        return tnn.GroupNorm(layer.num_groups, layer.num_channels, eps=layer.epsilon, affine=layer.affine)
    if isinstance(layer, MaxPool2DLayer):
        return tnn.MaxPool2d(layer.pool, layer.stride, layer.pad)
    if isinstance(layer, AvgPool2DLayer):
        return tnn.AvgPool2d(layer.pool, layer.stride, layer.pad, count_include_pad=True)   # divisor is always kh*kw
    if isinstance(layer, DropoutLayer):
        return tnn.Dropout(layer.dropout_rate)
    if isinstance(layer, FlattenLayer):
        return Flatten4D()
    if isinstance(layer, ActivationLayer):
        return _activation(layer.activation)
    if isinstance(layer, ResidualBlock):
        main = tnn.Sequential(*[_convert(l) for l in layer.main_path])
        sc = tnn.Sequential(*[_convert(l) for l in layer.shortcut_path]) if layer.shortcut_path else tnn.Identity()
        return Residual(main, sc, _activation(layer.final_activation) if layer.final_activation is not None
                        else tnn.Identity())
    raise ValueError(f"no torch counterpart for layer type {type(layer).__name__}")


def _pairs(layer, module) -> List[tuple]:
    """(our layer, torch module) for every parameterised leaf, in matching order."""
    if isinstance(layer, ResidualBlock):
        out = []
        for l, m in zip(layer.main_path, module.main):
            out += _pairs(l, m)
        if layer.shortcut_path:
            for l, m in zip(layer.shortcut_path, module.shortcut):
                out += _pairs(l, m)
        return out
    if isinstance(layer, (Conv2DLayer, DenseLayer, BatchNormLayer, GroupNormLayer)):
        return [(layer, module)]
    return []


def _copy(layer, module, to_torch: bool):
    def mv(ours, theirs):
        if ours is None or theirs is None:
            return
        if to_torch:
            theirs.data.copy_(ours.detach().reshape(theirs.shape))
        else:
            ours.copy_(theirs.detach().reshape(ours.shape))
    if isinstance(layer, (Conv2DLayer, DenseLayer)):
        mv(layer.weights, module.weight)
        if layer.use_bias:
            mv(layer.bias, module.bias)
    else:
        if layer.affine:
            mv(layer.gamma, module.weight)
            mv(layer.beta, module.bias)
        if isinstance(layer, BatchNormLayer):
            mv(layer.running_mean, module.running_mean)
            mv(layer.running_var, module.running_var)


def to_torch_module(model, copy_weights: bool = True) -> tnn.Sequential:
    """The equivalent ``torch.nn.Sequential`` on the model's device (parameters and BN statistics copied if the model
    is initialised and ``copy_weights``)."""
    layers = model.get_layers()
    net = tnn.Sequential(*[_convert(l) for l in layers])
    dev = model.get_device()
    net = net.to(dev if isinstance(dev, torch.device) else torch.device("cpu"))
    if copy_weights and layers and getattr(layers[0], "_initialized", False):
        with torch.no_grad():
            for l, m in zip(layers, net):
                for ll, mm in _pairs(l, m):
                    _copy(ll, mm, True)
    net.train(model.is_training())
    return net


def load_from_torch_module(model, net: tnn.Module):
    """Copy parameters and BN statistics of a module produced by ``to_torch_module`` (same architecture) back into
    ``model`` (which must be initialised)."""
    with torch.no_grad():
        for l, m in zip(model.get_layers(), net):
            for ll, mm in _pairs(l, m):
                _copy(ll, mm, False)
