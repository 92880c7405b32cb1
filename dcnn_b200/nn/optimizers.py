"""Optimizers (ref include/nn/optimizers.hpp:89-306, adam_kernels.cu:17-61, sgd_kernels.cu:17-35).

``attach(params, grads)`` then ``update()`` / ``clear_gradients()``.  On CUDA the update is ONE
multi-tensor kernel launch over a device-resident pointer table (the reference launches one kernel
per parameter tensor) and can zero the gradients in the same pass (``update(zero_grads=True)``).
Optimizer state can be exported/imported (``state_dict``) -- the reference never serialises it."""
from __future__ import annotations

from typing import List, Sequence

import torch

from .. import ops


class Optimizer:
    def __init__(self, learning_rate: float):
        self.learning_rate = float(learning_rate)
        self.params: List[torch.Tensor] = []
        self.grads: List[torch.Tensor] = []
        self._table = None  # backend-specific cached pointer table
        #: optional device tensor {lr, bias_corr1, bias_corr2}: step-varying scalars read from memory so
        #: a captured CUDA graph of the whole training step can be replayed (see train.TrainStep)
        self.hyper = None
        self._hyper_host = None

    def enable_device_hyper(self, device):
        import torch as _t
        self.hyper = _t.zeros(4, device=device, dtype=_t.float32)
        self._hyper_host = _t.zeros(4, dtype=_t.float32).pin_memory() if _t.cuda.is_available() else _t.zeros(4)

    def _hyper_values(self):
        return (self.learning_rate, 1.0, 1.0)

    def push_hyper(self):
        """Stage this step's scalars on the device (stream-ordered, outside any captured graph)."""
        if self.hyper is None:
            return
        import torch as _t
        if self.hyper.is_cuda and _t.cuda.is_current_stream_capturing():
            return
        lr, b1, b2 = self._hyper_values()
        self._hyper_host[0], self._hyper_host[1], self._hyper_host[2] = lr, b1, b2
        self.hyper.copy_(self._hyper_host, non_blocking=True)

    def attach(self, params: Sequence[torch.Tensor], grads: Sequence[torch.Tensor]):
        if len(params) != len(grads):
            raise ValueError("Parameters and gradients size mismatch")
        self.params, self.grads = list(params), list(grads)
        self._table = None
        self.on_attach()

    def on_attach(self):
        pass

    def advance(self):
        pass

    def update(self, zero_grads: bool = False):
        raise NotImplementedError

    def clear_gradients(self):
        for g in self.grads:
            g.zero_()

    def set_learning_rate(self, lr: float):
        self.learning_rate = float(lr)

    def get_learning_rate(self) -> float:
        return self.learning_rate

    def name(self) -> str:
        raise NotImplementedError

    def get_config(self) -> dict:
        raise NotImplementedError

    def clone(self):
        return OptimizerFactory.create_from_config(self.get_config())

    def state_dict(self) -> dict:
        return {}

    def load_state_dict(self, state: dict):
        pass


class SGD(Optimizer):
    """p -= lr*g, or with momentum: v = mu*v - lr*g; p += v."""

    def __init__(self, learning_rate: float = 0.01, momentum: float = 0.0):
        super().__init__(learning_rate)
        self.momentum = float(momentum)
        self.velocities: List[torch.Tensor] = []

    def on_attach(self):
        self.velocities = [torch.zeros_like(p) for p in self.params] if self.momentum > 0 else []

    def update(self, zero_grads: bool = False):
        if not self.params:
            return
        self.push_hyper()
        ops.sgd_update(self.params, self.grads, self.velocities, self.learning_rate, self.momentum,
                       zero_grads=zero_grads, hyper=self.hyper)

    def name(self):
        return "SGD"

    def get_config(self):
        return {"type": "sgd", "name": "SGD",
                "parameters": {"learning_rate": self.learning_rate, "momentum": self.momentum}}

    def state_dict(self):
        return {"velocities": self.velocities}

    def load_state_dict(self, state):
        for v, s in zip(self.velocities, state.get("velocities", [])):
            v.copy_(s)


class Adam(Optimizer):
    """Adam / AdamW with bias correction; L2 (``update += wd*lr*p``) or decoupled decay
    (``p -= wd*lr*p``) exactly as adam_kernels.cu:24-60."""

    def __init__(self, learning_rate: float = 0.001, beta1: float = 0.9, beta2: float = 0.999,
                 epsilon: float = 1e-8, weight_decay: float = 0.0, decouple_weight_decay: bool = False):
        super().__init__(learning_rate)
        self.beta1, self.beta2, self.epsilon = float(beta1), float(beta2), float(epsilon)
        self.weight_decay, self.decouple_weight_decay = float(weight_decay), bool(decouple_weight_decay)
        self.t = 0
        self.m: List[torch.Tensor] = []
        self.v: List[torch.Tensor] = []

    def on_attach(self):
        self.m = [torch.zeros_like(p) for p in self.params]
        self.v = [torch.zeros_like(p) for p in self.params]
        self.t = 0

    def bias_corrections(self):
        return 1.0 - self.beta1 ** self.t, 1.0 - self.beta2 ** self.t

    def _hyper_values(self):
        bc1, bc2 = self.bias_corrections()
        return (self.learning_rate, bc1, bc2)

    def advance(self):
        """Advance the step counter without launching (used when a captured graph is replayed)."""
        self.t += 1

    def update(self, zero_grads: bool = False):
        if not self.params:
            return
        self.t += 1
        bc1, bc2 = self.bias_corrections()
        self.push_hyper()
        ops.adam_update(self.params, self.grads, self.m, self.v, self.learning_rate, self.beta1,
                        self.beta2, self.epsilon, bc1, bc2, self.weight_decay,
                        self.decouple_weight_decay, zero_grads=zero_grads, hyper=self.hyper)

    def name(self):
        return "AdamW" if self.decouple_weight_decay else "Adam"

    def get_config(self):
        return {"type": "adamw" if self.decouple_weight_decay else "adam", "name": self.name(),
                "parameters": {"learning_rate": self.learning_rate, "beta1": self.beta1,
                               "beta2": self.beta2, "epsilon": self.epsilon,
                               "weight_decay": self.weight_decay,
                               "decouple_weight_decay": self.decouple_weight_decay}}

    def state_dict(self):
        return {"t": self.t, "m": self.m, "v": self.v}

    def load_state_dict(self, state):
        self.t = int(state.get("t", 0))
        for dst, src in zip(self.m, state.get("m", [])):
            dst.copy_(src)
        for dst, src in zip(self.v, state.get("v", [])):
            dst.copy_(src)


class OptimizerFactory:
    @staticmethod
    def create(name: str, learning_rate: float, momentum: float = 0.9) -> Optimizer:
        n = name.lower()
        if n == "sgd":
            return SGD(learning_rate, momentum)
        if n == "adam":
            return Adam(learning_rate)
        if n == "adamw":
            return Adam(learning_rate, decouple_weight_decay=True)
        raise ValueError(f"Unknown optimizer type: {name}")

    @staticmethod
    def create_from_config(config: dict) -> Optimizer:
        t = config.get("type", "").lower()
        p = config.get("parameters", {})
        lr = p.get("learning_rate", 0.001)
        if t == "sgd":
            return SGD(lr, p.get("momentum", 0.0))
        if t in ("adam", "adamw"):
            return Adam(lr, p.get("beta1", 0.9), p.get("beta2", 0.999), p.get("epsilon", 1e-8),
                        p.get("weight_decay", 0.0), p.get("decouple_weight_decay", t == "adamw"))
        raise ValueError(f"Unknown optimizer type: {config.get('type')}")
