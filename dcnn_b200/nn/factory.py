"""LayerFactory / LayerBuilder / free factory functions
(ref include/nn/layers.hpp:50-483).  JSON typing rule on load: integer -> size_t, float -> float,
bool, string (sequential.hpp:1104-1116) -- in Python the JSON types map directly."""
from __future__ import annotations

import json
from typing import Any, Callable, Dict, List, Sequence

from .activations import create_activation
from .layers import (ActivationLayer, AvgPool2DLayer, BatchNormLayer, Conv2DLayer, DenseLayer,
                     DropoutLayer, FlattenLayer, GroupNormLayer, Layer, MaxPool2DLayer)


def dense_layer(input_features, output_features, use_bias=True, name="dense"):
    return DenseLayer(input_features, output_features, use_bias, name)


def conv2d_layer(in_channels, out_channels, kernel_h, kernel_w, stride_h=1, stride_w=1, pad_h=0,
                 pad_w=0, use_bias=True, name="conv2d"):
    return Conv2DLayer(in_channels, out_channels, kernel_h, kernel_w, stride_h, stride_w, pad_h,
                       pad_w, use_bias, name)


def activation_layer(activation_name, name="activation"):
    return ActivationLayer(create_activation(activation_name), name)


def maxpool2d_layer(pool_h, pool_w, stride_h=0, stride_w=0, pad_h=0, pad_w=0, name="maxpool2d"):
    return MaxPool2DLayer(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w, name)


def avgpool2d_layer(pool_h, pool_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0, name="avgpool2d"):
    return AvgPool2DLayer(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w, name)


def dropout_layer(dropout_rate, name="dropout"):
    return DropoutLayer(dropout_rate, name)


def batchnorm_layer(num_features, epsilon=1e-5, momentum=0.1, affine=True, name="batchnorm"):
    return BatchNormLayer(num_features, epsilon, momentum, affine, name)


def groupnorm_layer(num_groups, num_channels, epsilon=1e-5, affine=True, name="groupnorm"):
    return GroupNormLayer(num_groups, num_channels, epsilon, affine, name)


def flatten_layer(name="flatten"):
    return FlattenLayer(name)


class LayerFactory:
    """String-keyed layer registry; ``create(type, {"name":..., "parameters": {...}})``."""

    _creators: Dict[str, Callable[[Dict[str, Any]], Layer]] = {}

    @classmethod
    def register_layer(cls, type_name, creator):
        cls._creators[type_name] = creator

    @classmethod
    def create(cls, type_name, config=None):
        if config is None:  # create(config) form
            config, type_name = type_name, type_name.get("type")
        if not cls._creators:
            cls.register_defaults()
        if type_name not in cls._creators:
            raise ValueError(f"Unknown layer type: {type_name}")
        return cls._creators[type_name](config)

    @classmethod
    def available_types(cls):
        if not cls._creators:
            cls.register_defaults()
        return list(cls._creators)

    @classmethod
    def register_defaults(cls):
        def P(cfg):
            return cfg.get("parameters", {})

        def nm(cfg):
            return cfg.get("name", "")

        cls._creators.update({
            "dense": lambda c: DenseLayer(P(c)["input_features"], P(c)["output_features"],
                                          P(c).get("use_bias", True), nm(c)),
            "conv2d": lambda c: Conv2DLayer(
                P(c)["in_channels"], P(c)["out_channels"], P(c)["kernel_h"], P(c)["kernel_w"],
                P(c).get("stride_h", 1), P(c).get("stride_w", 1), P(c).get("pad_h", 0),
                P(c).get("pad_w", 0), P(c).get("use_bias", True), nm(c)),
            "activation": lambda c: ActivationLayer(create_activation(P(c)["activation"]), nm(c)),
            "maxpool2d": lambda c: MaxPool2DLayer(
                P(c)["pool_h"], P(c)["pool_w"], P(c).get("stride_h", 0), P(c).get("stride_w", 0),
                P(c).get("pad_h", 0), P(c).get("pad_w", 0), nm(c)),
            "avgpool2d": lambda c: AvgPool2DLayer(
                P(c)["pool_h"], P(c)["pool_w"], P(c).get("stride_h", 0), P(c).get("stride_w", 0),
                P(c).get("pad_h", 0), P(c).get("pad_w", 0), nm(c)),
            "dropout": lambda c: DropoutLayer(P(c)["dropout_rate"], nm(c)),
            "batchnorm": lambda c: BatchNormLayer(P(c)["num_features"], P(c).get("epsilon", 1e-5),
                                                  P(c).get("momentum", 0.1), P(c).get("affine", True), nm(c)),
            "groupnorm": lambda c: GroupNormLayer(P(c)["num_groups"], P(c)["num_channels"],
                                                  P(c).get("epsilon", 1e-5), P(c).get("affine", True), nm(c)),
            "flatten": lambda c: FlattenLayer(nm(c)),
            "residual_block": cls._make_residual,
        })

    @classmethod
    def _make_residual(cls, c):
        from .blocks import ResidualBlock
        p = c.get("parameters", {})

        def parse(v):
            arr = json.loads(v) if isinstance(v, str) else (v or [])
            return [cls.create(sub.get("type", ""), {"name": sub.get("name", ""),
                                                     "parameters": sub.get("parameters", {})})
                    for sub in arr]
        main = parse(p.get("main_path", "[]"))
        sc = parse(p.get("shortcut_path", "[]")) if p.get("has_projection", False) else []
        return ResidualBlock(main, sc, p.get("activation", "relu"), c.get("name", ""))


class LayerBuilder:
    """Shape-inferring list builder (ref layers.hpp:297-483). Auto names use the running layer
    count, e.g. ``conv2d_0``, ``activation_2`` -- needed for JSON compatibility."""

    def __init__(self, name: str = "Block"):
        self.layers: List[Layer] = []
        self.input_shape: List[int] = []
        self.input_shape_set = False

    def input(self, shape: Sequence[int]):
        self.input_shape = [int(s) for s in shape]
        self.input_shape_set = True
        return self

    def is_input_shape_set(self):
        return self.input_shape_set

    def get_input_shape(self):
        return self.input_shape

    def get_current_shape(self):
        if not self.input_shape_set:
            raise RuntimeError("Input shape must be set before adding layers. Use .input() method first.")
        s = [1] + list(self.input_shape)
        for l in self.layers:
            s = l.compute_output_shape(s)
        return s

    def get_feature_count(self):
        s = self.get_current_shape()
        n = 1
        for d in s[1:]:
            n *= d
        return n

    def _n(self, prefix, name):
        return name if name else f"{prefix}_{len(self.layers)}"

    def dense(self, output_features, use_bias=True, name=""):
        self.layers.append(DenseLayer(self.get_feature_count(), output_features, use_bias,
                                      self._n("dense", name)))
        return self

    def conv2d(self, out_channels, kernel_h, kernel_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0,
               use_bias=True, name=""):
        s = self.get_current_shape()
        if len(s) < 4:
            raise RuntimeError("Conv2D requires 4D input (batch, channels, height, width).")
        self.layers.append(Conv2DLayer(s[1], out_channels, kernel_h, kernel_w, stride_h, stride_w,
                                       pad_h, pad_w, use_bias, self._n("conv2d", name)))
        return self

    def batchnorm(self, epsilon=1e-5, momentum=0.1, affine=True, name=""):
        s = self.get_current_shape()
        self.layers.append(BatchNormLayer(s[1], epsilon, momentum, affine, self._n("batchnorm", name)))
        return self

    def groupnorm(self, num_groups, epsilon=1e-5, affine=True, name=""):
        s = self.get_current_shape()
        self.layers.append(GroupNormLayer(int(num_groups), s[1], epsilon, affine,
                                          self._n("groupnorm", name)))
        return self

    def activation(self, activation_name, name=""):
        self.layers.append(activation_layer(activation_name, self._n("activation", name)))
        return self

    def maxpool2d(self, pool_h, pool_w, stride_h=0, stride_w=0, pad_h=0, pad_w=0, name=""):
        self.layers.append(MaxPool2DLayer(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w,
                                          self._n("maxpool2d", name)))
        return self

    def avgpool2d(self, pool_h, pool_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0, name=""):
        self.layers.append(AvgPool2DLayer(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w,
                                          self._n("avgpool2d", name)))
        return self

    def dropout(self, dropout_rate, name=""):
        self.layers.append(DropoutLayer(dropout_rate, self._n("dropout", name)))
        return self

    def flatten(self, name=""):
        self.layers.append(FlattenLayer(self._n("flatten", name)))
        return self

    def add_layer(self, layer):
        self.layers.append(layer)
        return self

    def build(self):
        if not self.input_shape_set:
            raise RuntimeError("Input shape must be set before building block. Use .input() method.")
        out, self.layers = self.layers, []
        return out
