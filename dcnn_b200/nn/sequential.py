"""Sequential container, Partition and SequentialBuilder
(ref include/nn/sequential.hpp:39-1340).

API parity: add/insert/remove/clear/layer_size, set_training, set_device, initialize, set_seed,
forward/backward(micro_batch_id), parameters/gradients (whole or by Partition),
compute_output_shape, forward/backward_complexity vectors, split, clone, get_config /
load_from_config / save_config / load_from_config_file, save_to_file / from_file /
load_weights_file (byte-compatible ``.json`` + ``.bin``), profiling summary.

Differences (deliberate, B200-first): no end-of-pass device synchronisation and no clone of the
result (sequential.hpp:489-500 does both); per-layer timing uses CUDA events on the compute
stream instead of host wall-clock around asynchronous launches (§5.1 of SURVEY); conv/BN/ReLU
chains execute as fused units (nn/fusion.py).
"""
from __future__ import annotations

import json
import os
import time
from typing import Dict, List, Optional, Sequence

import torch

from ..utils import runtime as _rt

from ..io.tensor_codec import read_tensor, write_tensor
from .blocks import ResidualBlock
from .factory import LayerBuilder, LayerFactory
from .fusion import build_units
from .layers import Layer


class Partition:
    """[start_layer, end_layer) (ref sequential.hpp:30-37)."""

    def __init__(self, start_layer: int, end_layer: int):
        self.start_layer, self.end_layer = int(start_layer), int(end_layer)

    def __repr__(self):
        return f"Partition({self.start_layer}, {self.end_layer})"

    def __eq__(self, o):
        return (self.start_layer, self.end_layer) == (o.start_layer, o.end_layer)


class Sequential:
    def __init__(self, name: str = "sequential"):
        self.name_ = name
        self.layers: List[Layer] = []
        self.training = True
        self.device = torch.device("cpu")
        self.enable_profiling_ = False
        self.after_unit_backward = None       # optional callable(unit_index), called after each unit's backward
        self.forward_times_us: Dict[str, float] = {}
        self.backward_times_us: Dict[str, float] = {}
        self._pending_events: List[tuple] = []
        self.fuse = True
        self._units = None
        #: compute dL/dinput in backward() (the reference always does and ships it back to the
        #: coordinator, which discards it -- coordinator.hpp:316-321; trainers turn this off)
        self.compute_input_gradient = True

    # -- container ---------------------------------------------------------------------------
    def add(self, layer: Layer):
        if layer is None:
            raise ValueError("Cannot add null layer")
        layer.set_training(self.training)
        layer.enable_profiling(self.enable_profiling_)
        self.layers.append(layer)
        self._units = None

    def insert(self, index: int, layer: Layer):
        if index > len(self.layers):
            raise IndexError("Insert index out of range")
        layer.set_training(self.training)
        self.layers.insert(index, layer)
        self._units = None

    def remove(self, index: int):
        if index >= len(self.layers):
            raise IndexError("Remove index out of range")
        del self.layers[index]
        self._units = None

    def clear(self):
        self.layers.clear()
        self._units = None
        self.clear_profiling_data()

    def layer_size(self) -> int:
        return len(self.layers)

    def __len__(self):
        return len(self.layers)

    def get_layers(self) -> List[Layer]:
        return list(self.layers)

    def name(self) -> str:
        return self.name_

    def set_name(self, name: str):
        self.name_ = name

    # -- state -------------------------------------------------------------------------------
    def set_seed(self, seed: int):
        for i, l in enumerate(self.layers):
            l.set_seed(int(seed) + 104729 * i)

    def initialize(self):
        for l in self.layers:
            l.initialize()

    def set_training(self, training: bool):
        self.training = bool(training)
        for l in self.layers:
            l.set_training(training)

    def is_training(self) -> bool:
        return self.training

    def set_device(self, device):
        """Accepts a torch.device, 'CPU'/'GPU' (ref DeviceType), 'cuda:i', or 'GPU:i'."""
        if isinstance(device, str):
            d = device.strip()
            up = d.upper()
            if up == "CPU" or up.startswith("CPU:"):
                device = torch.device("cpu")
            elif up == "GPU":
                device = torch.device("cuda", torch.cuda.current_device() if torch.cuda.is_available() else 0)
            elif up.startswith("GPU:"):
                device = torch.device("cuda", int(d.split(":")[1]))
            else:
                device = torch.device(d)
        self.device = torch.device(device)
        for l in self.layers:
            l.set_device(self.device)

    def get_device(self):
        return self.device

    # -- profiling ---------------------------------------------------------------------------
    def enable_profiling(self, enable: bool = True):
        self.enable_profiling_ = enable
        for l in self.layers:
            l.enable_profiling(enable)
        if not enable:
            self.clear_profiling_data()

    def is_profiling_enabled(self):
        return self.enable_profiling_

    def clear_profiling_data(self):
        self.forward_times_us.clear()
        self.backward_times_us.clear()
        self._pending_events.clear()

    def clear_forward_times(self):
        self.forward_times_us.clear()

    def clear_backward_times(self):
        self.backward_times_us.clear()

    def _resolve_events(self):
        if not self._pending_events:
            return
        torch.cuda.synchronize()
        for table, name, e0, e1 in self._pending_events:
            table[name] = table.get(name, 0.0) + e0.elapsed_time(e1) * 1000.0
        self._pending_events.clear()

    def get_forward_times(self):
        self._resolve_events()
        return dict(self.forward_times_us)

    def get_backward_times(self):
        self._resolve_events()
        return dict(self.backward_times_us)

    def _timed(self, table, name, fn, *a, **k):
        rec = _rt.ACTIVE_TRACE
        if rec is not None:        # Chrome-trace recording (utils/trace.py)
            return rec.timed(self.device, name, "forward" if table is self.forward_times_us else "backward",
                             lambda: self._timed_inner(table, name, fn, *a, **k))
        if _rt.NVTX and self.device.type == "cuda":   # DCNN_NVTX=1: one range per fusion unit and direction
            tag = ("fwd " if table is self.forward_times_us else "bwd ") + name
            with _rt.nvtx_range(tag):
                return self._timed_inner(table, name, fn, *a, **k)
        return self._timed_inner(table, name, fn, *a, **k)

    def _timed_inner(self, table, name, fn, *a, **k):
        if not self.enable_profiling_:
            return fn(*a, **k)
        if self.device.type == "cuda":
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*a, **k)
            e1.record()
            self._pending_events.append((table, name, e0, e1))
            return out
        t0 = time.perf_counter()
        out = fn(*a, **k)
        table[name] = table.get(name, 0.0) + (time.perf_counter() - t0) * 1e6
        return out

    def print_profiling_summary(self):
        """Same table as the reference (sequential.hpp:323-418): Forward/Backward/Total ms."""
        fw, bw = self.get_forward_times(), self.get_backward_times()
        if not fw and not bw:
            print("No profiling data available.")
            return
        line = "=" * 90
        print(line)
        print(f"Performance Profile: {self.name_}")
        print(line)
        print(f"{'Layer':<44}{'Forward (ms)':>15}{'Backward (ms)':>15}{'Total (ms)':>15}")
        print("-" * 90)
        tf = tb = 0.0
        for n in list(dict.fromkeys(list(fw) + list(bw))):
            f, b = fw.get(n, 0.0) / 1000.0, bw.get(n, 0.0) / 1000.0
            tf, tb = tf + f, tb + b
            print(f"{n[:43]:<44}{f:>15.3f}{b:>15.3f}{f + b:>15.3f}")
        print("-" * 90)
        print(f"{'TOTAL':<44}{tf:>15.3f}{tb:>15.3f}{tf + tb:>15.3f}")
        print(line)

    def print_layers_profiling_info(self):
        for l in self.layers:
            l.print_profiling_info()

    def print_cache_memory_summary(self):
        total = 0
        print(f"Cache memory summary: {self.name_}")
        for l in self.layers:
            b = l.cached_memory_bytes()
            total += b
            print(f"  {l.name:<40}{b / 1048576.0:>12.3f} MB")
        print(f"  {'TOTAL':<40}{total / 1048576.0:>12.3f} MB")

    # -- compute -----------------------------------------------------------------------------
    def _get_units(self):
        if self._units is None:
            for l in self.layers:
                if isinstance(l, ResidualBlock):
                    l.fuse = self.fuse
                    l._units_main = None
            self._units = build_units(self.layers, self.fuse)
        return self._units

    def _foreign_device(self) -> bool:
        """True when this model lives on a CUDA device that is not the calling thread's current device (kernels are
        launched on the current device's stream, so the call has to switch first)."""
        d = self.device
        return d.type == "cuda" and d.index is not None and torch.cuda.current_device() != d.index

    def forward(self, x: torch.Tensor, micro_batch_id: int = 0) -> torch.Tensor:
        if not self.layers:
            raise RuntimeError("Cannot forward through empty sequential model")
        if self._foreign_device():
            with torch.cuda.device(self.device):
                return self.forward(x, micro_batch_id)
        if x.device != self.device:
            x = x.to(self.device, non_blocking=True)
        for u in self._get_units():
            x = self._timed(self.forward_times_us, u.name, u.forward, x, micro_batch_id)
        return x

    def __call__(self, x, micro_batch_id: int = 0):
        return self.forward(x, micro_batch_id)

    def backward(self, grad: torch.Tensor, micro_batch_id: int = 0) -> Optional[torch.Tensor]:
        if not self.layers:
            raise RuntimeError("Cannot backward through empty sequential model")
        if self._foreign_device():
            with torch.cuda.device(self.device):
                return self.backward(grad, micro_batch_id)
        if grad.device != self.device:
            grad = grad.to(self.device, non_blocking=True)
        units = self._get_units()
        first = units[0].layers[0]
        if hasattr(first, "need_input_grad"):
            first.need_input_grad = self.compute_input_gradient
        hook = self.after_unit_backward
        for idx in range(len(units) - 1, -1, -1):
            u = units[idx]
            grad, _ = self._timed(self.backward_times_us, u.name, u.backward, grad, micro_batch_id)
            if hook is not None:
                hook(idx)          # e.g. DataParallel: the gradients of units >= idx are final for this pass
            if grad is None:
                break
        return grad

    # -- parameters --------------------------------------------------------------------------
    def _slice(self, part: Optional[Partition]):
        if part is None:
            return self.layers
        if part.start_layer >= len(self.layers) or part.end_layer > len(self.layers) \
                or part.start_layer >= part.end_layer:
            raise IndexError("Partition indices out of range")
        return self.layers[part.start_layer:part.end_layer]

    def parameters(self, part: Optional[Partition] = None) -> List[torch.Tensor]:
        return [p for l in self._slice(part) for p in l.parameters()]

    def gradients(self, part: Optional[Partition] = None) -> List[torch.Tensor]:
        return [g for l in self._slice(part) for g in l.gradients()]

    def clear_gradients(self):
        for g in self.gradients():
            g.zero_()

    def clear_cache(self):
        for l in self.layers:
            l.clear_cache()

    def load_parameters(self, parameters: Sequence[torch.Tensor]):
        own = self.parameters()
        if len(own) != len(parameters):
            raise RuntimeError(f"Parameter count mismatch: expected {len(own)} but got {len(parameters)}")
        for i, (p, q) in enumerate(zip(own, parameters)):
            if tuple(p.shape) != tuple(q.shape):
                raise RuntimeError(f"Parameter shape mismatch at index {i}")
            p.copy_(q.to(p.device))

    def buffers(self) -> Dict[str, torch.Tensor]:
        out = {}
        for i, l in enumerate(self.layers):
            for k, v in l.buffers().items():
                out[f"{i}.{k}"] = v
        return out

    # -- shapes / complexity -----------------------------------------------------------------
    def compute_output_shape(self, input_shape: Sequence[int]) -> List[int]:
        s = list(input_shape)
        for l in self.layers:
            s = l.compute_output_shape(s)
        return s

    def _per_layer(self, fn, input_shape):
        out, s = [], list(input_shape)
        for l in self.layers:
            out.append(getattr(l, fn)(s))
            s = l.compute_output_shape(s)
        return out

    def forward_complexity(self, input_shape):
        return self._per_layer("forward_complexity", input_shape)

    def backward_complexity(self, input_shape):
        return self._per_layer("backward_complexity", input_shape)

    def forward_flops(self, input_shape):
        return self._per_layer("forward_flops", input_shape)

    def backward_flops(self, input_shape):
        return self._per_layer("backward_flops", input_shape)

    def print_summary(self, input_shape: Sequence[int]):
        print("=" * 100)
        print(f"Model Summary: {self.name_}")
        print("=" * 100)
        print(f"{'Layer (Type)':<36}{'Input Shape':<22}{'Output Shape':<22}{'Fwd FLOPs':>10}{'Bwd FLOPs':>10}")
        s = list(input_shape)
        for l in self.layers:
            o = l.compute_output_shape(s)
            nm = f"{l.name or '<unnamed>'} ({l.type()})"
            print(f"{nm[:35]:<36}{str(tuple(s)):<22}{str(tuple(o)):<22}"
                  f"{l.forward_flops(s):>10.3g}{l.backward_flops(s):>10.3g}")
            s = o
        print("-" * 100)

    def print_config(self):
        print(json.dumps(self.get_config(), indent=2))

    # -- config / checkpoint -----------------------------------------------------------------
    def get_config(self, partition: Optional[Partition] = None) -> dict:
        layers = self._slice(partition)
        return {"name": self.name_, "is_training": self.training,
                "layers": [{"type": l.type(), "name": l.get_config()["name"],
                            "parameters": l.get_config()["parameters"]} for l in layers]}

    @staticmethod
    def load_from_config(config: dict) -> "Sequential":
        model = Sequential(config.get("name", "sequential"))
        model.training = config.get("is_training", True)
        for lj in config.get("layers", []):
            model.add(LayerFactory.create(lj["type"], {"name": lj.get("name", ""),
                                                       "parameters": lj.get("parameters", {})}))
        return model

    def save_config(self, filepath: str):
        with open(filepath, "w") as f:
            f.write(json.dumps(self.get_config(), indent=4))

    @staticmethod
    def load_from_config_file(filepath: str) -> "Sequential":
        with open(filepath) as f:
            return Sequential.load_from_config(json.load(f))

    def save_to_file(self, path: str, save_state: bool = False):
        """Writes ``path.json`` + ``path.bin`` exactly like the reference
        (sequential.hpp:832-861).  ``save_state`` additionally writes the additive
        ``path.state.bin/.json`` sidecar (BN running stats) that the reference loader ignores."""
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        with open(path + ".json", "w") as f:
            f.write(json.dumps(self.get_config(), indent=4))
        with open(path + ".bin", "wb") as f:
            for l in self.layers:
                if l.has_parameters():
                    for p in l.parameters():
                        write_tensor(f, p)
        if save_state:
            bufs = self.buffers()
            with open(path + ".state.bin", "wb") as f:
                for k in bufs:
                    write_tensor(f, bufs[k])
            with open(path + ".state.json", "w") as f:
                json.dump({"buffers": list(bufs)}, f, indent=4)

    def load_weights_file(self, path: str):
        with open(path, "rb") as f:
            for l in self.layers:
                l.initialize()
                if l.has_parameters():
                    for p in l.parameters():
                        t = read_tensor(f)
                        if t.numel() != p.numel():
                            raise RuntimeError(f"{l.name}: weight record has {t.numel()} elements, expected {p.numel()}")
                        p.copy_(t.reshape(p.shape).to(p.device))
        st = path[:-4] + ".state.bin" if path.endswith(".bin") else None
        if st and os.path.exists(st):
            with open(st, "rb") as f:
                for k, b in self.buffers().items():
                    b.copy_(read_tensor(f).reshape(b.shape).to(b.device))

    @staticmethod
    def from_file(path: str, device=None) -> "Sequential":
        with open(path + ".json") as f:
            model = Sequential.load_from_config(json.load(f))
        if device is not None:
            model.set_device(device)
        model.load_weights_file(path + ".bin")
        return model

    def clone(self) -> "Sequential":
        c = Sequential(self.name_)
        c.set_training(self.training)
        for l in self.layers:
            c.add(l.clone())
        return c

    def split(self, partitions: Sequence[Partition], copy_weights: bool = False) -> List["Sequential"]:
        """One ``Sequential`` per partition, named ``<name>_part_<i>`` (sequential.hpp:967-986).
        The reference clones configs only; ``copy_weights=True`` also carries the parameters and
        BN running statistics over (used by the equivalence tests and checkpoint scatter)."""
        if not partitions:
            raise ValueError("Partitions vector is empty")
        stages = []
        for part in partitions:
            if part.start_layer >= len(self.layers) or part.end_layer > len(self.layers) \
                    or part.start_layer >= part.end_layer:
                raise IndexError("Invalid partition range")
            st = Sequential(f"{self.name_}_part_{len(stages)}")
            st.fuse = self.fuse
            for i in range(part.start_layer, part.end_layer):
                st.add(self.layers[i].clone())
            if copy_weights:
                st.initialize()
                src = Sequential("tmp")
                src.layers = self.layers[part.start_layer:part.end_layer]
                st.load_parameters(src.parameters())
                for (k, b), (_, sb) in zip(st.buffers().items(), src.buffers().items()):
                    b.copy_(sb)
            stages.append(st)
        return stages


class SequentialBuilder:
    """Fluent builder with shape inference (ref sequential.hpp:1154-1340).

    Quirk kept for JSON compatibility: auto-generated names use
    ``model.layer_size()``, which is 0 until ``build()``, so every auto-named top-level layer is
    ``<type>_0``; BNs inside the residual helpers are all named ``bn0``."""

    def __init__(self, name: str = "sequential"):
        self.model = Sequential(name)
        self.lb = LayerBuilder()

    def get_current_shape(self):
        return self.lb.get_current_shape()

    def input(self, shape):
        self.lb.input(shape)
        return self

    def _n(self, prefix, name):
        return name if name else f"{prefix}_{self.model.layer_size()}"

    def dense(self, output_features, use_bias=True, name=""):
        self.lb.dense(output_features, use_bias, self._n("dense", name))
        return self

    def conv2d(self, out_channels, kernel_h, kernel_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0,
               use_bias=True, name=""):
        self.lb.conv2d(out_channels, kernel_h, kernel_w, stride_h, stride_w, pad_h, pad_w, use_bias,
                       self._n("conv2d", name))
        return self

    def batchnorm(self, epsilon=1e-5, momentum=0.1, affine=True, name=""):
        self.lb.batchnorm(epsilon, momentum, affine, self._n("batchnorm", name))
        return self

    def groupnorm(self, num_groups, epsilon=1e-5, affine=True, name=""):
        self.lb.groupnorm(num_groups, epsilon, affine, self._n("groupnorm", name))
        return self

    def activation(self, activation_name, name=""):
        self.lb.activation(activation_name, self._n("activation", name))
        return self

    def maxpool2d(self, pool_h, pool_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0, name=""):
        self.lb.maxpool2d(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w, self._n("maxpool2d", name))
        return self

    def avgpool2d(self, pool_h, pool_w, stride_h=1, stride_w=1, pad_h=0, pad_w=0, name=""):
        self.lb.avgpool2d(pool_h, pool_w, stride_h, stride_w, pad_h, pad_w, self._n("avgpool2d", name))
        return self

    def dropout(self, dropout_rate, name=""):
        self.lb.dropout(dropout_rate, self._n("dropout", name))
        return self

    def flatten(self, name=""):
        self.lb.flatten(self._n("flatten", name))
        return self

    def residual(self, main_path, shortcut=None, activation_name="relu", name=""):
        self.lb.add_layer(ResidualBlock(main_path, shortcut or [], activation_name,
                                        self._n("residual_block", name)))
        return self

    def basic_residual_block(self, in_channels, out_channels, stride=1, name="basic_residual_block"):
        cur = self.lb.get_current_shape()
        ishape = [in_channels, cur[2], cur[3]]
        main = (LayerBuilder().input(ishape)
                .conv2d(out_channels, 3, 3, stride, stride, 1, 1, True)
                .batchnorm(1e-5, 0.1, True, "bn0")
                .activation("relu")
                .conv2d(out_channels, 3, 3, 1, 1, 1, 1, True)
                .batchnorm(1e-5, 0.1, True, "bn0")
                .build())
        sc = []
        if stride != 1 or in_channels != out_channels:
            sc = (LayerBuilder().input(ishape)
                  .conv2d(out_channels, 1, 1, stride, stride, 0, 0, False)
                  .batchnorm(1e-5, 0.1, True, "bn0")
                  .build())
        self.lb.add_layer(ResidualBlock(main, sc, "relu", self._n("basic_residual_block", name)))
        return self

    def bottleneck_residual_block(self, in_channels, mid_channels, out_channels, stride=1,
                                  name="bottleneck_residual_block"):
        cur = self.lb.get_current_shape()
        ishape = [in_channels, cur[2], cur[3]]
        main = (LayerBuilder().input(ishape)
                .conv2d(mid_channels, 1, 1, 1, 1, 0, 0, False)
                .batchnorm(1e-3, 0.1, True, "bn0")
                .activation("relu")
                .conv2d(mid_channels, 3, 3, stride, stride, 1, 1, False)
                .batchnorm(1e-3, 0.1, True, "bn0")
                .activation("relu")
                .conv2d(out_channels, 1, 1, 1, 1, 0, 0, False)
                .batchnorm(1e-3, 0.1, True, "bn0")
                .build())
        sc = []
        if stride != 1 or in_channels != out_channels:
            sc = (LayerBuilder().input(ishape)
                  .conv2d(out_channels, 1, 1, stride, stride, 0, 0, False)
                  .batchnorm(1e-3, 0.1, True, "bn0")
                  .build())
        self.lb.add_layer(ResidualBlock(main, sc, "relu", name))
        return self

    def add_layer(self, layer):
        self.lb.add_layer(layer)
        return self

    def build(self) -> Sequential:
        if not self.lb.is_input_shape_set():
            raise RuntimeError("Input shape must be set before building model. Use .input() method.")
        for l in self.lb.build():
            self.model.add(l)
        return self.model
