"""Loss functions (ref include/nn/loss.hpp:41-472).  Targets are one-hot floats ``[N,classes,1,1]``;
the loss value is the mean over the batch and the gradient is pre-scaled by 1/batch.

B200 path: one fused kernel computes loss + gradient + correct-count in a single pass and leaves
the scalars on the device (no host synchronisation per call, unlike loss_ops.cu:379-419).  Use
``forward_backward`` on the hot path; ``compute_loss`` / ``compute_gradient`` keep the reference
API and synchronise only when the Python float is requested."""
from __future__ import annotations


import torch

from .. import ops


class Loss:
    kind = ""
    display = "Loss"
    param = 0.0

    def name(self) -> str:
        return self.display

    def num_parameters(self) -> int:
        """Tunable scalars of the loss (Huber's delta); 0 for the others (ref loss.hpp)."""
        return 1 if self.kind == "huber" else 0

    def forward_backward(self, predictions, targets, want_grad=True):
        """-> (loss scalar tensor on device, grad | None, correct-count tensor | None)."""
        return ops.loss_forward_backward(self.kind, predictions, targets, self.param, want_grad)

    def compute_loss(self, predictions, targets) -> float:
        loss, _, _ = self.forward_backward(predictions, targets, want_grad=False)
        return float(loss)

    def compute_gradient(self, predictions, targets) -> torch.Tensor:
        _, g, _ = self.forward_backward(predictions, targets, want_grad=True)
        return g

    def get_config(self):
        return {"type": self.kind, "name": self.display, "parameters": {}}

    def clone(self):
        return LossFactory.create_from_config(self.get_config())


class CrossEntropyLoss(Loss):
    """-log(clamp(p, eps, 1-eps)) on probabilities; grad (p - t)/B."""
    kind, display = "crossentropy", "CrossEntropyLoss"

    def __init__(self, epsilon: float = 1e-15):
        self.param = float(epsilon)

    def get_config(self):
        return {"type": self.kind, "name": self.display, "parameters": {"epsilon": self.param}}


class SoftmaxCrossEntropyLoss(Loss):
    kind, display = "softmax_crossentropy", "SoftmaxCrossEntropyLoss"


class LogSoftmaxCrossEntropyLoss(Loss):
    kind, display = "logsoftmax_crossentropy", "LogSoftmaxCrossEntropyLoss"


class MSELoss(Loss):
    kind, display = "mse", "MSELoss"


class MAELoss(Loss):
    kind, display = "mae", "MAELoss"


class HuberLoss(Loss):
    kind, display = "huber", "HuberLoss"

    def __init__(self, delta: float = 1.0):
        self.param = float(delta)

    def get_config(self):
        return {"type": self.kind, "name": self.display, "parameters": {"delta": self.param}}

    def set_delta(self, delta: float):
        self.param = float(delta)

    def get_delta(self) -> float:
        return self.param


class LossFactory:
    _ALIASES = {
        "crossentropy": "crossentropy", "ce": "crossentropy",
        "softmax_crossentropy": "softmax_crossentropy", "softmax_ce": "softmax_crossentropy",
        "logsoftmax_crossentropy": "logsoftmax_crossentropy", "logsoftmax_ce": "logsoftmax_crossentropy",
        "mse": "mse", "mean_squared_error": "mse", "mae": "mae", "mean_absolute_error": "mae",
        "huber": "huber"}

    @classmethod
    def create(cls, loss_type: str) -> Loss:
        return cls.create_from_config({"type": loss_type, "parameters": {}})

    @classmethod
    def create_from_config(cls, config) -> Loss:
        t = cls._ALIASES.get(config.get("type", ""))
        p = config.get("parameters", {})
        if t == "crossentropy":
            return CrossEntropyLoss(p.get("epsilon", 1e-15))
        if t == "softmax_crossentropy":
            return SoftmaxCrossEntropyLoss()
        if t == "logsoftmax_crossentropy":
            return LogSoftmaxCrossEntropyLoss()
        if t == "mse":
            return MSELoss()
        if t == "mae":
            return MAELoss()
        if t == "huber":
            return HuberLoss(p.get("delta", 1.0))
        raise ValueError(f"Unknown loss type: {config.get('type')}")

    create_crossentropy = staticmethod(lambda epsilon=1e-15: CrossEntropyLoss(epsilon))
    create_softmax_crossentropy = staticmethod(lambda: SoftmaxCrossEntropyLoss())
    create_logsoftmax_crossentropy = staticmethod(lambda: LogSoftmaxCrossEntropyLoss())
    create_mse = staticmethod(lambda: MSELoss())
    create_mae = staticmethod(lambda: MAELoss())
    create_huber = staticmethod(lambda delta=1.0: HuberLoss(delta))
