"""Activation functions + string-keyed factory (ref include/nn/activations.hpp:27-64,
include/nn/activations_impl/*).  ``apply`` works in place like the reference; ``forward`` /
``backward`` are the out-of-place forms the layers use."""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch

from .. import ops


class ActivationFunction:
    kind = "linear"
    alpha = 0.0

    def name(self) -> str:
        return self.kind

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.activation_forward(x, self.kind, self.alpha)

    def backward(self, pre_activation: torch.Tensor, grad: torch.Tensor) -> torch.Tensor:
        return ops.activation_backward(pre_activation, grad, self.kind, self.alpha)

    # reference-style in-place API
    def apply(self, x: torch.Tensor) -> torch.Tensor:
        x.copy_(self.forward(x))
        return x

    def compute_gradient_inplace(self, pre_activation: torch.Tensor, grad: torch.Tensor) -> torch.Tensor:
        grad.copy_(self.backward(pre_activation, grad))
        return grad

    def clone(self):
        return type(self)() if self.alpha == 0.0 else type(self)(self.alpha)


class ReLU(ActivationFunction):
    kind = "relu"


class LeakyReLU(ActivationFunction):
    kind = "leaky_relu"

    def __init__(self, negative_slope: float = 0.01):
        self.alpha = float(negative_slope)


class ELU(ActivationFunction):
    kind = "elu"

    def __init__(self, alpha: float = 1.0):
        self.alpha = float(alpha)


class Sigmoid(ActivationFunction):
    kind = "sigmoid"


class Tanh(ActivationFunction):
    kind = "tanh"


class Softmax(ActivationFunction):
    kind = "softmax"


class Linear(ActivationFunction):
    kind = "linear"


class ActivationFactory:
    _creators: Dict[str, Callable[[], Optional[ActivationFunction]]] = {}

    @classmethod
    def register_activation(cls, name, creator):
        cls._creators[name] = creator

    @classmethod
    def register_defaults(cls):
        cls._creators.update({
            "none": lambda: None, "relu": ReLU, "leaky_relu": lambda: LeakyReLU(0.01),
            "sigmoid": Sigmoid, "softmax": Softmax, "linear": Linear, "tanh": Tanh, "elu": ELU})

    @classmethod
    def create(cls, name: str):
        if not cls._creators:
            cls.register_defaults()
        if name not in cls._creators:
            raise ValueError(f"Unknown activation function: {name}")
        return cls._creators[name]()

    @classmethod
    def get_available_activations(cls):
        if not cls._creators:
            cls.register_defaults()
        return list(cls._creators)


def create_activation(name: str):
    return ActivationFactory.create(name)
