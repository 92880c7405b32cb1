"""Training drivers (ref include/nn/train.hpp:30-460): ``TrainingConfig`` (same .env keys and defaults),
``train_class_epoch`` / ``validate_class_model`` / ``train_classification_model`` and the regression twins.

B200-first differences: the per-batch work is a :class:`TrainStep` -- forward, fused loss(+grad+accuracy),
backward, one multi-tensor optimizer launch that also zeroes the gradients -- with *no* host synchronisation
inside the step (the reference syncs >= 4 times per step).  ``TrainStep`` can capture the whole
step in a CUDA graph; loss/accuracy scalars stay on the device and are read back every
``PROGRESS_PRINT_INTERVAL`` batches and at the end of the epoch."""
from __future__ import annotations

import os
import time
from dataclasses import dataclass
from typing import List, Optional

import torch

from . import ops
from .nn.loss import Loss
from .nn.optimizers import Optimizer
from .nn.sequential import Sequential
from .utils import runtime as _runtime
from .utils.env import get_env
from .utils.runtime import checkpoint_on_signal, metrics_from_env, watchdog_from_env


@dataclass
class TrainingConfig:
    epochs: int = 10
    batch_size: int = 32
    lr_decay_factor: float = 0.9
    lr_decay_interval: int = 5
    progress_print_interval: int = 100
    profiler_type: str = "NONE"      # NONE | NORMAL | CUMULATIVE
    num_threads: int = 8
    print_layer_profiling: bool = False
    num_microbatches: int = 2
    device_type: str = "CPU"         # CPU | GPU
    use_cuda_graph: bool = False
    snapshot_dir: str = "model_snapshots"

    def load_from_env(self):
        self.epochs = get_env("EPOCHS", self.epochs)
        self.batch_size = get_env("BATCH_SIZE", self.batch_size)
        self.lr_decay_factor = get_env("LR_DECAY_FACTOR", self.lr_decay_factor)
        self.lr_decay_interval = get_env("LR_DECAY_INTERVAL", self.lr_decay_interval)
        self.progress_print_interval = get_env("PROGRESS_PRINT_INTERVAL", self.progress_print_interval)
        p = get_env("PROFILER_TYPE", "NONE")
        self.profiler_type = p if p in ("NORMAL", "CUMULATIVE") else "NONE"
        self.num_threads = get_env("NUM_THREADS", self.num_threads)
        self.print_layer_profiling = get_env("PRINT_LAYER_PROFILING", False)
        self.num_microbatches = get_env("NUM_MICROBATCHES", self.num_microbatches)
        self.device_type = "CPU" if get_env("DEVICE_TYPE", "CPU") == "CPU" else "GPU"
        self.use_cuda_graph = get_env("USE_CUDA_GRAPH", self.use_cuda_graph)
        return self

    def print_config(self):
        print("Training Configuration:")
        print(f"  Epochs: {self.epochs}")
        print(f"  Batch Size: {self.batch_size}")
        print(f"  LR Decay Factor: {self.lr_decay_factor}")
        print(f"  LR Decay Interval (epochs): {self.lr_decay_interval}")
        print(f"  Progress Print Interval (batches): {self.progress_print_interval}")
        print(f"  Number of Threads: {self.num_threads}")
        print(f"  Profiler Type: {self.profiler_type.capitalize()}")
        print(f"  Print Layer Profiling Info: {'Yes' if self.print_layer_profiling else 'No'}")
        print(f"  Number of Microbatches: {self.num_microbatches}")
        print(f"  Device Type: {self.device_type}")

    def torch_device(self):
        if self.device_type == "GPU":
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device("cpu")


def split_microbatches(t: torch.Tensor, m: int) -> List[torch.Tensor]:
    """B/M each, the last microbatch takes the remainder (ref tensor/cpu/tensor_ops.hpp:675-689)."""
    b = t.shape[0]
    if m <= 1 or b < m:
        return [t]
    size = b // m
    out = [t[i * size:(i + 1) * size] for i in range(m - 1)]
    out.append(t[(m - 1) * size:])
    return out


class TrainStep:
    """One optimisation step on one device: forward -> loss+grad+accuracy -> backward -> update(+zero grads).

    ``num_microbatches > 1`` runs the microbatches back to back with gradient accumulation (each microbatch
    gradient is scaled by 1/microbatch_size and the loss is the mean over microbatches, which is the
    reference pipeline's arithmetic, coordinator.hpp:323-325).  ``use_cuda_graph`` captures the whole step after
    two eager warm-up steps; inputs must then keep a fixed shape."""

    def __init__(self, model: Sequential, optimizer: Optimizer, loss_fn: Loss, use_cuda_graph: bool = False,
                 num_microbatches: int = 1, data_parallel=None):
        self.model, self.optimizer, self.loss_fn = model, optimizer, loss_fn
        self.num_microbatches = max(1, int(num_microbatches))
        self.dp = data_parallel
        self.use_cuda_graph = bool(use_cuda_graph) and model.get_device().type == "cuda"
        self._graph = None
        self._static = None
        self._calls = 0
        self._trace_path = os.environ.get("DCNN_TRACE_FILE", "")
        self._trace_total = self._trace_left = int(os.environ.get("DCNN_TRACE_STEPS", "3") or 3) if self._trace_path else 0
        self._trace = None
        if self.use_cuda_graph:
            optimizer.enable_device_hyper(model.get_device())

    def _eager(self, x, t):
        xs = split_microbatches(x, self.num_microbatches)
        ts = split_microbatches(t, self.num_microbatches)
        loss_acc = correct_acc = None
        grads = []
        for i, (xi, ti) in enumerate(zip(xs, ts)):
            y = self.model.forward(xi, i)
            loss, g, correct = self.loss_fn.forward_backward(y, ti)
            grads.append(g)
            loss_acc = loss if loss_acc is None else loss_acc + loss
            if correct is not None:
                correct_acc = correct if correct_acc is None else correct_acc + correct
        for i in reversed(range(len(xs))):
            if self.dp is not None:
                self.dp.final_backward = (i == 0)     # bucketed DP may start reducing finished buckets during this pass
            self.model.backward(grads[i], i)
        if self.dp is not None:
            self.dp.reduce_and_update(self.optimizer)
        else:
            self.optimizer.update(zero_grads=True)
        if len(xs) > 1:
            loss_acc = loss_acc / len(xs)
        return loss_acc, correct_acc

    def __call__(self, x: torch.Tensor, t: torch.Tensor):
        _runtime.kick()                   # feeds any running StepWatchdog (STEP_TIMEOUT_S)
        if self._trace_path and self._trace_left > 0:
            return self._traced_call(x, t)
        return self._call(x, t)

    def _traced_call(self, x, t):
        """DCNN_TRACE_FILE: Chrome trace of the first DCNN_TRACE_STEPS steps (eager steps only -- a replayed CUDA graph
        does not pass through the per-unit hooks)."""
        from .utils.trace import TraceRecorder
        if self._trace is None:
            root, ext = os.path.splitext(self._trace_path)
            rank = int(os.environ.get("RANK", "0"))
            self._trace = TraceRecorder(self._trace_path if rank == 0 else f"{root}.rank{rank}{ext or '.json'}")
            self._trace.__enter__()
        self._trace.mark(f"step {self._calls if self.use_cuda_graph else self._trace_total - self._trace_left}")
        try:
            return self._call(x, t)
        finally:
            self._trace_left -= 1
            if self._trace_left == 0 or self._graph is not None:
                self._trace_left = 0
                self._trace.__exit__(None, None, None)
                print(f"[dcnn] trace written to {self._trace.path}")

    def _call(self, x: torch.Tensor, t: torch.Tensor):
        if not self.use_cuda_graph:
            return self._eager(x, t)
        self._calls += 1
        if self._graph is None:
            if self._calls <= 2:  # eager warm-up: lazy init, cudaFuncSetAttribute, pointer tables
                return self._eager(x, t)
            self._static = (x.clone(), t.clone())
            g = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            # host-side state that the captured (not executed) step advances: rolled back if the capture fails
            saved_t = getattr(self.optimizer, "t", None)
            try:
                with torch.cuda.graph(g):
                    out = self._eager(*self._static)
            except Exception as e:      # keep training: a failed capture costs launch overhead, not the run
                import sys
                print(f"[dcnn] CUDA graph capture failed ({type(e).__name__}: {e}); continuing without graphs",
                      file=sys.stderr)
                self.use_cuda_graph, self._static = False, None
                if saved_t is not None:
                    self.optimizer.t = saved_t
                if self.dp is not None:
                    self.dp.reset_step_bookkeeping()
                torch.cuda.synchronize()
                return self._eager(x, t)
            self._graph, self._out = g, out
            # the capture itself did not execute: fall through to replay with this batch
            if hasattr(self.optimizer, "t"):
                self.optimizer.t -= 1
        self._static[0].copy_(x, non_blocking=True)
        self._static[1].copy_(t, non_blocking=True)
        self.optimizer.advance()
        self.optimizer.push_hyper()
        self._graph.replay()
        ops.weights_changed()             # the replayed update wrote the parameters behind torch's back
        return self._out


def compute_class_corrects(predictions: torch.Tensor, targets: torch.Tensor) -> int:
    """argmax-vs-one-hot correct count (ref utils/utils_extended.hpp:11-41)."""
    b = predictions.shape[0]
    return int((predictions.reshape(b, -1).argmax(1) == targets.reshape(b, -1).argmax(1).to(predictions.device)).sum())


def compute_class_accuracy(predictions, targets) -> float:
    return compute_class_corrects(predictions, targets) / max(1, predictions.shape[0])


def get_memory_usage_kb() -> int:
    try:
        with open("/proc/self/status") as f:
            for line in f:
                if line.startswith("VmRSS:"):
                    return int(line.split()[1])
    except OSError:
        pass
    return 0


def _to_device(t, device):
    if t.device == device:
        return t
    if device.type == "cuda" and not t.is_pinned():
        t = t.pin_memory()
    return t.to(device, non_blocking=True)


def train_class_epoch(model, train_loader, optimizer, loss_function, config: Optional[TrainingConfig] = None,
                      step: Optional[TrainStep] = None):
    config = config or TrainingConfig()
    print("Starting training epoch...")
    model.set_training(True)
    train_loader.shuffle()
    train_loader.reset()
    step = step or TrainStep(model, optimizer, loss_function, config.use_cuda_graph)
    device = model.get_device()
    total_loss = torch.zeros((), device=device, dtype=torch.float64)
    total_corrects = torch.zeros((), device=device, dtype=torch.float64)
    num_samples = num_batches = 0
    print(f"Training batches: {train_loader.num_batches()}")
    while True:
        batch = train_loader.get_next_batch()
        if batch is None:
            break
        data, labels = batch
        num_batches += 1
        num_samples += data.shape[0]
        loss, corrects = step(_to_device(data, device), _to_device(labels, device))
        total_loss += loss.double()
        if corrects is not None:
            total_corrects += corrects.double()
        if num_batches % config.progress_print_interval == 0:
            if model.is_profiling_enabled():
                if config.print_layer_profiling:
                    model.print_layers_profiling_info()
                model.print_profiling_summary()
            print(f"Batch ID: {num_batches}, Batch's Loss: {float(loss):.4f}, Cumulative Accuracy: "
                  f"{float(total_corrects) * 100.0 / num_samples:.2f}%")
        if model.is_profiling_enabled() and config.profiler_type == "NORMAL":
            model.clear_profiling_data()
    print()
    return float(total_loss) / max(1, num_batches), float(total_corrects) / max(1, num_samples)


def validate_class_model(model, test_loader, loss_function):
    model.set_training(False)
    test_loader.reset()
    device = model.get_device()
    val_loss = torch.zeros((), device=device, dtype=torch.float64)
    val_corrects = torch.zeros((), device=device, dtype=torch.float64)
    batches = samples = 0
    print("Starting validation...")
    while True:
        batch = test_loader.get_next_batch()
        if batch is None:
            break
        data, labels = batch
        pred = model.forward(_to_device(data, device))
        loss, _, corrects = loss_function.forward_backward(pred, _to_device(labels, device), want_grad=False)
        val_loss += loss.double()
        if corrects is not None:
            val_corrects += corrects.double()
        batches += 1
        samples += data.shape[0]
    model.set_training(True)
    return float(val_loss) / max(1, batches), float(val_corrects) / max(1, samples)


def _mean_over_replicas(values, group, device):
    """Average a list of Python floats over the replicas of ``group`` (one small all-reduce)."""
    import torch.distributed as dist
    t = torch.tensor(values, dtype=torch.float64, device=device if device.type == "cuda" else "cpu")
    dist.all_reduce(t, group=group)
    return (t / dist.get_world_size(group)).tolist()


def train_classification_model(model, train_loader, test_loader, optimizer, loss_function,
                               config: Optional[TrainingConfig] = None, data_parallel=None):
    """``data_parallel``: a :class:`dcnn_b200.parallel.dp.DataParallel` built on this model/optimizer (one process per
    replica).  Every replica then trains on its shard of the data (``config.batch_size`` is the per-replica batch), the
    reported losses / accuracies are averages over replicas, and only replica 0 prints and writes snapshots.  The
    reference has no data parallelism; without ``data_parallel`` this is the reference's loop."""
    config = config or TrainingConfig()
    model.compute_input_gradient = False
    dp = data_parallel
    main = dp is None or dp.rank == 0
    if dp is None:
        optimizer.attach(model.parameters(), model.gradients())
    else:
        for ld in (train_loader, test_loader):
            ld.set_seed(1234)                 # identical shuffles on every replica, disjoint shards
            ld.set_shard(dp.rank, dp.world)
    if not main:
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):      # replicas other than 0 stay quiet
            return _train_classification(model, train_loader, test_loader, optimizer, loss_function, config, dp, main)
    return _train_classification(model, train_loader, test_loader, optimizer, loss_function, config, dp, main)


def _train_classification(model, train_loader, test_loader, optimizer, loss_function, config, dp, main):
    # RESUME_FROM=<path written by CHECKPOINT_EVERY / checkpoint_on_signal>: continue where that run stopped -- weights,
    # BatchNorm statistics, optimizer moments and step, learning rate, epoch counter, RNG and data-order state
    resume = get_env("RESUME_FROM", "")
    state = None
    if resume:
        from .io import load_training_state
        state = load_training_state(resume, model, optimizer)
        ls = state.get("extra", {}).get("loaders")
        if ls:
            train_loader.set_rng_state(ls[0])
            test_loader.set_rng_state(ls[1])
        print(f"Resumed from {resume}: epoch {state['epoch']}, lr {optimizer.get_learning_rate():.6g}")
    # PREFETCH_DEPTH=k: stream the training set instead of materialising every batch -- batches are built k ahead on a
    # worker thread (fresh augmentation every epoch, sample-level reshuffle) and copied to the GPU on a side stream
    depth = get_env("PREFETCH_DEPTH", 0)
    if depth > 0:
        from .data import PrefetchLoader
        dev = model.get_device()
        train_loader = PrefetchLoader(train_loader, config.batch_size, depth,
                                      dev if getattr(dev, "type", "cpu") == "cuda" else None)
    train_loader.prepare_batches(config.batch_size)
    test_loader.prepare_batches(config.batch_size)
    model.enable_profiling(config.profiler_type in ("NORMAL", "CUMULATIVE"))
    print(f"Training batches: {train_loader.num_batches()}")
    print(f"Validation batches: {test_loader.num_batches()}")
    c, h, w = train_loader.get_image_shape()
    model.print_summary([config.batch_size, c, h, w])
    if config.device_type == "CPU":
        torch.set_num_threads(max(1, config.num_threads))
    step = TrainStep(model, optimizer, loss_function, config.use_cuda_graph, data_parallel=dp)
    best_val_accuracy = 0.0
    history = []
    metrics = metrics_from_env()
    epoch_box = [0]
    ckpt = os.path.join(config.snapshot_dir, model.name() + ".interrupted")
    with watchdog_from_env("train step") as wd, checkpoint_on_signal(ckpt, model, optimizer, lambda: epoch_box[0]):
        history = _classification_epochs(model, train_loader, test_loader, optimizer, loss_function, config, step,
                                         metrics, wd, epoch_box, dp, main, state)
    return history


def _classification_epochs(model, train_loader, test_loader, optimizer, loss_function, config, step, metrics, wd,
                           epoch_box, dp=None, main=True, resumed=None):
    best_val_accuracy = 0.0
    history = []
    start_epoch = 0
    if resumed is not None:
        start_epoch = int(resumed.get("epoch", 0))
        best_val_accuracy = float(resumed.get("extra", {}).get("best_val_accuracy", 0.0))
    every = get_env("CHECKPOINT_EVERY", 0)          # epochs between resumable checkpoints (0 = off)
    for epoch in range(start_epoch, config.epochs):
        epoch_box[0] = epoch
        print(f"Epoch {epoch + 1}/{config.epochs}")
        t0 = time.perf_counter()
        tl, ta = train_class_epoch(model, train_loader, optimizer, loss_function, config, step)
        if model.get_device().type == "cuda":
            torch.cuda.synchronize()
        train_ms = (time.perf_counter() - t0) * 1000
        vl, va = validate_class_model(model, test_loader, loss_function)
        if dp is not None:      # equal shard sizes: the mean of the replicas' means is the global mean
            tl, ta, vl, va = _mean_over_replicas([tl, ta, vl, va], dp.group, model.get_device())
        if va > best_val_accuracy and not main:
            best_val_accuracy = va
        elif va > best_val_accuracy:
            best_val_accuracy = va
            print(f"New best validation accuracy: {best_val_accuracy * 100:.2f}%")
            try:
                path = os.path.join(config.snapshot_dir, model.name())
                model.save_to_file(path, save_state=True)
                print(f"Model saved to {path}")
            except OSError as e:
                print(f"Error saving model: {e}")
        print("-" * 60)
        print(f"Epoch {epoch + 1}/{config.epochs} completed in {train_ms:.0f}ms")
        print(f"Training   - Loss: {tl:.4f}, Accuracy: {ta * 100:.2f}%")
        print(f"Validation - Loss: {vl:.4f}, Accuracy: {va * 100:.2f}%")
        print("=" * 60)
        history.append({"epoch": epoch + 1, "train_loss": tl, "train_acc": ta, "val_loss": vl, "val_acc": va,
                        "train_ms": train_ms})
        n_img = train_loader.num_batches() * config.batch_size * (dp.world if dp is not None else 1)
        metrics.log(kind="epoch", images_per_s=n_img / max(train_ms, 1e-9) * 1000.0, rss_kb=get_memory_usage_kb(),
                    lr=optimizer.get_learning_rate(), **history[-1])
        wd.kick()
        if model.is_profiling_enabled():
            model.clear_profiling_data()
        if (epoch + 1) % config.lr_decay_interval == 0:
            cur = optimizer.get_learning_rate()
            optimizer.set_learning_rate(cur * config.lr_decay_factor)
            print(f"Learning rate decayed: {cur:.6f} -> {cur * config.lr_decay_factor:.6f}")
        if every > 0 and (epoch + 1) % every == 0 and main:
            # written BEFORE the next epoch's batches are drawn, so that a resumed run re-draws exactly the same batches
            from .io import save_training_state
            path = os.path.join(config.snapshot_dir, model.name() + ".checkpoint")
            save_training_state(path, model, optimizer, epoch=epoch + 1,
                                extra={"best_val_accuracy": best_val_accuracy,
                                       "loaders": [train_loader.get_rng_state(), test_loader.get_rng_state()]})
            print(f"Checkpoint written to {path} (resume with RESUME_FROM={path})")
        train_loader.prepare_batches(config.batch_size)  # re-apply augmentation
        test_loader.prepare_batches(config.batch_size)
        print(f"{get_memory_usage_kb() // 1024} MB of memory used.")
    return history


# ------------------------------------------------------------------------------------------ regression
def train_reg_epoch(model, train_loader, optimizer, loss_function, config=None, step=None):
    config = config or TrainingConfig()
    model.set_training(True)
    train_loader.shuffle()
    train_loader.reset()
    step = step or TrainStep(model, optimizer, loss_function, config.use_cuda_graph)
    device = model.get_device()
    total = torch.zeros((), device=device, dtype=torch.float64)
    n = 0
    while True:
        batch = train_loader.get_next_batch()
        if batch is None:
            break
        data, targets = batch
        loss, _ = step(_to_device(data, device), _to_device(targets, device))
        total += loss.double()
        n += 1
        if n % config.progress_print_interval == 0:
            print(f"Batch ID: {n}, Batch's Loss: {float(loss):.6f}")
    return float(total) / max(1, n)


def validate_reg_model(model, test_loader, loss_function):
    model.set_training(False)
    test_loader.reset()
    device = model.get_device()
    total = torch.zeros((), device=device, dtype=torch.float64)
    err = torch.zeros((), device=device, dtype=torch.float64)
    n = 0
    while True:
        batch = test_loader.get_next_batch()
        if batch is None:
            break
        data, targets = batch
        pred = model.forward(_to_device(data, device))
        tgt = _to_device(targets, device)
        loss, _, _ = loss_function.forward_backward(pred, tgt, want_grad=False)
        total += loss.double()
        err += (pred.reshape(tgt.shape[0], -1) - tgt.reshape(tgt.shape[0], -1)).abs().mean().double()
        n += 1
    model.set_training(True)
    return float(total) / max(1, n), float(err) / max(1, n)


def train_regression_model(model, train_loader, test_loader, optimizer, loss_function, config=None):
    config = config or TrainingConfig()
    model.compute_input_gradient = False
    optimizer.attach(model.parameters(), model.gradients())
    train_loader.prepare_batches(config.batch_size)
    test_loader.prepare_batches(config.batch_size)
    step = TrainStep(model, optimizer, loss_function, config.use_cuda_graph)
    best = float("inf")
    history = []
    metrics = metrics_from_env()
    epoch_box = [0]
    ckpt = os.path.join(config.snapshot_dir, model.name() + ".interrupted")
    with watchdog_from_env("train step"), checkpoint_on_signal(ckpt, model, optimizer, lambda: epoch_box[0]):
        return _regression_epochs(model, train_loader, test_loader, optimizer, loss_function, config, step, best, history,
                                  metrics, epoch_box)


def _regression_epochs(model, train_loader, test_loader, optimizer, loss_function, config, step, best, history, metrics,
                       epoch_box):
    for epoch in range(config.epochs):
        epoch_box[0] = epoch
        print(f"Epoch {epoch + 1}/{config.epochs}")
        t0 = time.perf_counter()
        tl = train_reg_epoch(model, train_loader, optimizer, loss_function, config, step)
        ms = (time.perf_counter() - t0) * 1000
        vl, ve = validate_reg_model(model, test_loader, loss_function)
        if vl < best:
            best = vl
            print(f"New best validation loss: {best:.6f}")
            try:
                model.save_to_file(os.path.join(config.snapshot_dir, model.name()), save_state=True)
            except OSError as e:
                print(f"Error saving model: {e}")
        print(f"Epoch {epoch + 1}/{config.epochs} completed in {ms:.0f}ms")
        print(f"Training   - Loss: {tl:.6f}")
        print(f"Validation - Loss: {vl:.6f}, Mean abs error: {ve:.6f}")
        history.append({"epoch": epoch + 1, "train_loss": tl, "val_loss": vl, "val_err": ve, "train_ms": ms})
        metrics.log(kind="epoch", lr=optimizer.get_learning_rate(), rss_kb=get_memory_usage_kb(), **history[-1])
        if (epoch + 1) % config.lr_decay_interval == 0:
            optimizer.set_learning_rate(optimizer.get_learning_rate() * config.lr_decay_factor)
        train_loader.prepare_batches(config.batch_size)
        test_loader.prepare_batches(config.batch_size)
    return history
