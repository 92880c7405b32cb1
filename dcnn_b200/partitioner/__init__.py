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
