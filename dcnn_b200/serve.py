"""Inference path: BatchNorm folding and a fixed-shape, graph-replayed predictor.

The reference evaluates a trained network with the training graph in eval mode (examples/mnist_cnn_test.cpp,
include/nn/train.hpp:172-215): every BatchNorm stays a separate layer and every batch pays the launch sequence again.
For serving:

  * ``fold_batchnorm(model)`` returns a copy in which every Conv2D -> BatchNorm pair (also inside residual blocks) is
    one Conv2D with bias: ``w' = w * gamma / sqrt(var + eps)``, ``b' = beta + (b - mean) * gamma / sqrt(var + eps)``.
    The BatchNorm kernels disappear from the inference graph; the result is numerically the eval-mode network.
  * ``Predictor`` owns a static input buffer of a fixed batch size, runs the (folded) network once eagerly and then --
    on CUDA -- replays the whole forward pass from one CUDA graph.  Smaller requests are padded, larger ones are cut
    into full batches; ``predict`` returns logits, ``predict_proba`` / ``predict_classes`` the usual reductions.
    ``stream_batches`` overlaps the host-to-device copy of request i+1 with the forward pass of request i
    (pinned staging, side stream)."""
from __future__ import annotations

from typing import Iterable, Iterator, List, Sequence

import torch

from .nn.blocks import ResidualBlock
from .nn.layers import BatchNormLayer, Conv2DLayer, Layer
from .nn.sequential import Sequential


def _fold_pair(conv: Conv2DLayer, bn: BatchNormLayer) -> Conv2DLayer:
    out = Conv2DLayer(conv.in_channels, conv.out_channels, conv.kernel[0], conv.kernel[1], conv.stride[0],
                      conv.stride[1], conv.pad[0], conv.pad[1], True, conv.name)
    out.set_device(conv.device)
    invstd = torch.rsqrt(bn.running_var.double() + bn.epsilon)
    scale = invstd * (bn.gamma.double().reshape(-1) if bn.affine else 1.0)
    shift = bn.beta.double().reshape(-1) if bn.affine else torch.zeros_like(scale)
    w = conv.weights.double() * scale.reshape(-1, 1, 1, 1)
    b0 = conv.bias.double().reshape(-1) if conv.use_bias else torch.zeros_like(scale)
    b = shift + (b0 - bn.running_mean.double()) * scale
    out.weights = out._to_dev(w.float(), True)
    out.bias = out._to_dev(b.float().reshape(-1, 1, 1, 1))
    out.weight_grad = torch.zeros_like(out.weights)
    out.bias_grad = torch.zeros_like(out.bias)
    out._initialized = True
    return out


def _copy_layer(layer: Layer) -> Layer:
    """Structural clone with the parameters and buffers copied."""
    new = layer.clone()
    new.set_device(layer.device)
    new.initialize()
    for dst, src in zip(new.parameters(), layer.parameters()):
        dst.copy_(src)
    for k, v in layer.buffers().items():
        new.buffers()[k].copy_(v)
    return new


def _fold_list(layers: Sequence[Layer]) -> List[Layer]:
    out: List[Layer] = []
    i = 0
    while i < len(layers):
        l = layers[i]
        nxt = layers[i + 1] if i + 1 < len(layers) else None
        if isinstance(l, Conv2DLayer) and isinstance(nxt, BatchNormLayer) and nxt.num_features == l.out_channels:
            out.append(_fold_pair(l, nxt))
            i += 2
        elif isinstance(l, ResidualBlock):
            out.append(ResidualBlock(_fold_list(l.main_path), _fold_list(l.shortcut_path), l.activation_type, l.name))
            out[-1].set_device(l.device)
            out[-1]._initialized = True
            i += 1
        else:
            out.append(_copy_layer(l))
            i += 1
    return out


def fold_batchnorm(model: Sequential) -> Sequential:
    """Eval-mode copy of ``model`` with every Conv2D -> BatchNorm pair folded into one biased convolution."""
    if not model.layers or not getattr(model.layers[0], "_initialized", False):
        raise RuntimeError("fold_batchnorm needs an initialised (trained or loaded) model")
    with torch.no_grad():
        folded = Sequential(model.name() + "_folded")
        for l in _fold_list(model.get_layers()):
            folded.add(l)
    folded.set_device(model.get_device())
    folded.set_training(False)
    return folded


def count_layers(model: Sequential, kind) -> int:
    def walk(layers):
        n = 0
        for l in layers:
            if isinstance(l, ResidualBlock):
                n += walk(l.main_path) + walk(l.shortcut_path)
            elif isinstance(l, kind):
                n += 1
        return n
    return walk(model.get_layers())


class Predictor:
    def __init__(self, model: Sequential, input_shape: Sequence[int], batch_size: int = 64, fold_bn: bool = True,
                 use_cuda_graph: bool = True):
        """``input_shape`` is one sample's (C, H, W).  The model is copied (folded or not) and left untouched."""
        self.batch_size = int(batch_size)
        self.input_shape = tuple(int(d) for d in input_shape)
        src = model
        with torch.no_grad():
            if fold_bn:
                self.model = fold_batchnorm(src)
            else:
                self.model = Sequential(src.name() + "_eval")
                for l in src.get_layers():
                    self.model.add(_copy_layer(l))
                self.model.set_device(src.get_device())
        self.model.set_training(False)
        self.device = self.model.get_device()
        self._cuda = self.device.type == "cuda"
        self._x = torch.zeros((self.batch_size,) + self.input_shape, device=self.device)
        self._graph = None
        self._y = None
        self._want_graph = bool(use_cuda_graph) and self._cuda
        self._stream = torch.cuda.Stream(self.device) if self._cuda else None
        self.requests = 0

    # -- one full static batch ------------------------------------------------------------------------------------
    def _run_static(self) -> torch.Tensor:
        if not self._want_graph:
            return self.model.forward(self._x)
        if self._graph is None:
            for _ in range(2):                  # lazy initialisation, kernel attribute set-up, descriptor caches
                self.model.forward(self._x)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g):
                    self._y = self.model.forward(self._x)
                self._graph = g
            except Exception as e:
                import sys
                print(f"[dcnn] CUDA graph capture of the predictor failed ({type(e).__name__}: {e}); running eagerly",
                      file=sys.stderr)
                self._want_graph = False
                torch.cuda.synchronize(self.device)
                return self.model.forward(self._x)
        self._graph.replay()
        return self._y

    def _check(self, x: torch.Tensor):
        if x.dim() != 4 or tuple(x.shape[1:]) != self.input_shape:
            raise ValueError(f"Predictor expects [N, {', '.join(map(str, self.input_shape))}] inputs, got {tuple(x.shape)}")

    @torch.no_grad()
    def predict(self, x: torch.Tensor) -> torch.Tensor:
        """Logits [N, classes, 1, 1] on the predictor's device for any N >= 0."""
        self._check(x)
        self.requests += 1
        outs = []
        for s in range(0, x.shape[0], self.batch_size):
            chunk = x[s:s + self.batch_size]
            n = chunk.shape[0]
            self._x[:n].copy_(chunk, non_blocking=True)
            if n < self.batch_size:
                self._x[n:].zero_()             # padding rows: defined values, results discarded
            outs.append(self._run_static()[:n].clone())
        if not outs:
            probe = self._run_static()
            return probe[:0].clone()
        return outs[0] if len(outs) == 1 else torch.cat(outs)

    def predict_proba(self, x: torch.Tensor) -> torch.Tensor:
        y = self.predict(x)
        return torch.softmax(y.reshape(y.shape[0], -1), dim=1)

    def predict_classes(self, x: torch.Tensor) -> torch.Tensor:
        y = self.predict(x)
        return y.reshape(y.shape[0], -1).argmax(dim=1)

    @torch.no_grad()
    def stream_batches(self, batches: Iterable[torch.Tensor]) -> Iterator[torch.Tensor]:
        """Logits for a stream of host batches (each at most ``batch_size`` samples).  On CUDA the copy of the next
        batch into a second device buffer runs on a side stream while the current batch is computed."""
        if not self._cuda:
            for b in batches:
                yield self.predict(b)
            return
        it = iter(batches)
        staged = [torch.zeros_like(self._x) for _ in range(2)]
        ready = [torch.cuda.Event() for _ in range(2)]
        freed = [torch.cuda.Event() for _ in range(2)]
        main = torch.cuda.current_stream(self.device)

        def stage(slot, b):
            self._check(b)
            n = b.shape[0]
            with torch.cuda.stream(self._stream):
                self._stream.wait_event(freed[slot])
                src = b if b.is_pinned() or b.is_cuda else b.pin_memory()
                staged[slot][:n].copy_(src, non_blocking=True)
                if n < self.batch_size:
                    staged[slot][n:].zero_()
                ready[slot].record(self._stream)
            return n

        for e in freed:
            e.record(main)
        nxt = next(it, None)
        slot = 0
        n = stage(slot, nxt) if nxt is not None else 0
        while nxt is not None:
            cur_slot, cur_n = slot, n
            nxt = next(it, None)
            if nxt is not None:
                slot ^= 1
                n = stage(slot, nxt)
            main.wait_event(ready[cur_slot])
            self._x.copy_(staged[cur_slot], non_blocking=True)
            freed[cur_slot].record(main)
            self.requests += 1
            yield self._run_static()[:cur_n].clone()

    def summary(self) -> str:
        return (f"Predictor({self.model.name()}, batch {self.batch_size}, device {self.device}, "
                f"cuda_graph={self._graph is not None}, conv {count_layers(self.model, Conv2DLayer)}, "
                f"batchnorm {count_layers(self.model, BatchNormLayer)})")
