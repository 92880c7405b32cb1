"""Model -> pipeline-stage partitioners (ref include/partitioner/partitioner.hpp:6-13,
naive_partitioner.hpp:7-34).  ``NaivePartitioner`` is bit-compatible with the reference (equal
*count* of top-level layers, remainder to the first partitions).  ``BalancedPartitioner`` is the
piece the reference lists as TODO ("Fix and finish the partitioner", TODO:2): contiguous split that
minimises the maximum per-stage cost using the layers' forward+backward complexity models."""
from __future__ import annotations

from typing import List, Optional, Sequence

from ..nn.sequential import Partition


class Partitioner:
    def get_partitions(self, layers: Sequence, num_partitions: int) -> List[Partition]:
        raise NotImplementedError


class NaivePartitioner(Partitioner):
    def get_partitions(self, layers, num_partitions):
        if num_partitions < 1:
            raise ValueError("Number of partitions must be at least 1")
        total = len(layers)
        base, rem = divmod(total, num_partitions)
        parts, start = [], 0
        for i in range(num_partitions):
            end = start + base + (1 if i < rem else 0)
            parts.append(Partition(start, end))
            start = end
        return parts


class BalancedPartitioner(Partitioner):
    """Min-max contiguous partition by cost (dynamic programming, O(S*L^2), L is tens of layers).

    cost(layer) = forward_complexity + backward_complexity at ``input_shape`` (batch included), or
    user-provided measured per-layer times (``costs=``, e.g. from Sequential profiling)."""

    def __init__(self, input_shape: Optional[Sequence[int]] = None, costs: Optional[Sequence[float]] = None):
        self.input_shape = list(input_shape) if input_shape is not None else None
        self.costs = list(costs) if costs is not None else None

    def layer_costs(self, layers):
        if self.costs is not None:
            if len(self.costs) != len(layers):
                raise ValueError("costs length must equal number of layers")
            return [float(c) for c in self.costs]
        if self.input_shape is None:
            raise ValueError("BalancedPartitioner needs input_shape or costs")
        out, s = [], list(self.input_shape)
        for l in layers:
            out.append(float(l.forward_flops(s) + l.backward_flops(s)) + 1.0)
            s = l.compute_output_shape(s)
        return out

    def get_partitions(self, layers, num_partitions):
        n = len(layers)
        if num_partitions < 1:
            raise ValueError("Number of partitions must be at least 1")
        if num_partitions > n:
            raise ValueError("More partitions than layers")
        c = self.layer_costs(layers)
        pre = [0.0]
        for v in c:
            pre.append(pre[-1] + v)
        INF = float("inf")
        best = [[INF] * (n + 1) for _ in range(num_partitions + 1)]
        cut = [[0] * (n + 1) for _ in range(num_partitions + 1)]
        best[0][0] = 0.0
        for k in range(1, num_partitions + 1):
            for j in range(k, n - (num_partitions - k) + 1):
                for i in range(k - 1, j):
                    v = max(best[k - 1][i], pre[j] - pre[i])
                    if v < best[k][j]:
                        best[k][j], cut[k][j] = v, i
        parts, j = [], n
        for k in range(num_partitions, 0, -1):
            i = cut[k][j]
            parts.append(Partition(i, j))
            j = i
        return parts[::-1]


def measure_layer_costs(model, input_shape: Sequence[int], warmup: int = 1, iters: int = 2) -> List[float]:
    """Measured forward+backward time (ms) per top-level layer of an initialised ``model`` on its device, for
    ``BalancedPartitioner(costs=...)``: what the reference's "load balancing (coming soon)" note asks for.

    One synthetic batch of ``input_shape`` is pushed through the whole model with the Sequential's own per-unit timers
    (CUDA events on a GPU, wall clock on the CPU).  A fusion unit (Conv -> BN -> ReLU ...) is timed as one piece; its time
    is spread over its layers so that the partitioner sees a cost for every layer."""
    import torch
    was_profiling, was_training = model.is_profiling_enabled(), model.is_training()
    model.set_training(True)
    model.enable_profiling(True)
    dev = model.get_device()
    x = torch.randn(*input_shape, device=dev)
    acc = {}
    saved_buffers = {k: v.clone() for k, v in model.buffers().items()}   # BatchNorm running statistics
    try:
        for it in range(warmup + iters):
            model.clear_profiling_data()
            y = model.forward(x, 0)
            model.backward(torch.ones_like(y) / max(1, y.numel()), 0)
            if dev.type == "cuda":
                torch.cuda.synchronize(dev)
            if it >= warmup:
                f, b = model.get_forward_times(), model.get_backward_times()
                for k in set(f) | set(b):
                    acc[k] = acc.get(k, 0.0) + (f.get(k, 0.0) + b.get(k, 0.0)) / 1000.0 / iters
    finally:
        model.clear_profiling_data()
        model.enable_profiling(was_profiling)
        model.set_training(was_training)
        model.clear_gradients()
        model.clear_cache()
        for k, v in model.buffers().items():
            v.copy_(saved_buffers[k])
    costs = []
    for u in model._get_units():
        t = acc.get(u.name, 0.0)
        costs.extend([max(t, 1e-6) / len(u.layers)] * len(u.layers))
    if len(costs) != model.layer_size():
        raise RuntimeError("fusion units do not cover the model's layers")
    return costs
