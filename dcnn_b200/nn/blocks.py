"""ResidualBlock: ``out = act(F(x) + shortcut(x))`` (ref include/nn/blocks_impl/residual_block.hpp:30-475).

Config compat: sub-layers are serialised as JSON *strings* under ``main_path`` / ``shortcut_path``
(residual_block.hpp:371-440).  Parameter order is all main-path layers, then shortcut layers
(residual_block.hpp:240-250).

Execution differs from the reference on purpose: when the main path ends in BatchNorm, the
residual add and the final ReLU are folded into that BN's apply pass, and in backward the ReLU
mask is folded into the BN reduce/apply passes while the two branch gradients are summed by the
dgrad epilogue (beta=1) instead of a separate add kernel.
"""
from __future__ import annotations

import json
from typing import Dict, List, Optional

import torch

from .. import ops
from .activations import create_activation
from .fusion import ConvBNActUnit, build_units
from .layers import Layer, _cache_bytes


class ResidualBlock(Layer):
    type_name = "residual_block"

    def __init__(self, main_path: List[Layer], shortcut_path: Optional[List[Layer]] = None,
                 activation: str = "relu", name: str = "residual_block"):
        super().__init__(name)
        self.main_path = list(main_path)
        self.shortcut_path = list(shortcut_path or [])
        self.activation_type = activation
        self.final_activation = create_activation(activation)
        self._pre_act: Dict[int, torch.Tensor] = {}
        self._units_main = self._units_sc = None
        self.fuse = True

    # -- plumbing ---------------------------------------------------------------------------
    def _all(self):
        return self.main_path + self.shortcut_path

    def initialize(self):
        for i, l in enumerate(self._all()):
            if self.seed is not None:
                l.set_seed(self.seed + 7919 * (i + 1))
            l.initialize()
        self._initialized = True

    def set_training(self, training):
        super().set_training(training)
        for l in self._all():
            l.set_training(training)

    def set_device(self, device):
        self.device = torch.device(device)
        for l in self._all():
            l.set_device(device)

    def _build(self):
        if self._units_main is None:
            self._units_main = build_units(self.main_path, self.fuse)
            self._units_sc = build_units(self.shortcut_path, self.fuse)

    def _fusable_tail(self):
        u = self._units_main[-1] if self._units_main else None
        return (isinstance(u, ConvBNActUnit) and u.can_fuse_residual()
                and self.activation_type in ("relu", "none", "linear"))

    # -- compute ----------------------------------------------------------------------------
    def forward(self, x, micro_batch_id=0):
        self._build()
        mb = micro_batch_id
        sc = x
        for u in self._units_sc:
            sc = u.forward(sc, mb)
        h = x
        if self._fusable_tail():
            for u in self._units_main[:-1]:
                h = u.forward(h, mb)
            return self._units_main[-1].forward(h, mb, residual=sc,
                                                relu=self.activation_type == "relu")
        for u in self._units_main:
            h = u.forward(h, mb)
        pre = ops.add(h, sc)
        if self.final_activation is None:
            return pre
        if self.training:
            self._pre_act[mb] = pre
        return self.final_activation.forward(pre)

    def backward(self, grad, micro_batch_id=0):
        self._build()
        mb = micro_batch_id
        if self._fusable_tail():
            g_main, g_branch = self._units_main[-1].backward(grad, mb, want_masked=True)
            rest = self._units_main[:-1]
        else:
            g_branch = grad
            if self.final_activation is not None:
                g_branch = self.final_activation.backward(self._pre_act[mb], grad)
            g_main, rest = g_branch, self._units_main
        identity = not self._units_sc
        # an identity shortcut's gradient is g_branch itself; let the last dgrad accumulate into it
        # when it is a private buffer (i.e. produced by the fused ReLU-masked BN backward).
        private = g_branch is not grad
        for k, u in enumerate(reversed(rest)):
            last = k == len(rest) - 1
            acc = g_branch if (last and identity and private) else None
            g_main, _ = u.backward(g_main, mb, accumulate_into=acc)
        if identity:
            if not rest or not private:
                g_main = ops.add(g_main, g_branch)
            return g_main
        g_sc = g_branch
        for k, u in enumerate(reversed(self._units_sc)):
            last = k == len(self._units_sc) - 1
            g_sc, _ = u.backward(g_sc, mb, accumulate_into=g_main if last else None)
        return g_sc

    # -- parameters -------------------------------------------------------------------------
    def parameters(self):
        return [p for l in self._all() for p in l.parameters()]

    def gradients(self):
        return [g for l in self._all() for g in l.gradients()]

    def buffers(self):
        out = {}
        for i, l in enumerate(self.main_path):
            for k, v in l.buffers().items():
                out[f"main.{i}.{k}"] = v
        for i, l in enumerate(self.shortcut_path):
            for k, v in l.buffers().items():
                out[f"shortcut.{i}.{k}"] = v
        return out

    def has_parameters(self):
        return any(l.has_parameters() for l in self._all())

    def rebind_parameters(self, params, grads):
        used = 0
        for l in self._all():
            used += l.rebind_parameters(params[used:], grads[used:])
        return used

    def clear_gradients(self):
        for l in self._all():
            l.clear_gradients()

    # -- metadata ---------------------------------------------------------------------------
    def get_config(self):
        def ser(layers):
            return json.dumps([{"type": l.type(), "name": l.get_config()["name"],
                                "parameters": l.get_config()["parameters"]} for l in layers])
        return {"name": self.name, "parameters": {
            "activation": self.activation_type, "has_projection": bool(self.shortcut_path),
            "main_path": ser(self.main_path), "shortcut_path": ser(self.shortcut_path)}}

    def compute_output_shape(self, s):
        for l in self.main_path:
            s = l.compute_output_shape(s)
        return list(s)

    def _sum(self, fn, s):
        """Main path + shortcut path + the element-wise addition over the block's output (forward) / the gradient
        summation (backward), as the reference counts it (residual_block.hpp:313-359)."""
        total, cur = 0, list(s)
        for l in self.main_path:
            total += getattr(l, fn)(cur)
            cur = l.compute_output_shape(cur)
        add = 1
        for d in cur:
            add *= int(d)
        cur = list(s)
        for l in self.shortcut_path:
            total += getattr(l, fn)(cur)
            cur = l.compute_output_shape(cur)
        return total + add

    def forward_flops(self, s):
        return self._sum("forward_flops", s)

    def backward_flops(self, s):
        return self._sum("backward_flops", s)

    def forward_complexity(self, s):
        return self._sum("forward_complexity", s)

    def backward_complexity(self, s):
        return self._sum("backward_complexity", s)

    def cached_memory_bytes(self):
        return _cache_bytes(self._pre_act) + sum(l.cached_memory_bytes() for l in self._all())

    def clear_cache(self):
        self._pre_act.clear()
        for l in self._all():
            l.clear_cache()

    def get_main_path(self):
        return self.main_path

    def get_shortcut_path(self):
        return self.shortcut_path


def residual_block(main_path, shortcut_path=None, activation="relu", name="residual_block"):
    return ResidualBlock(main_path, shortcut_path, activation, name)
