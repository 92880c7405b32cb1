"""Pipeline training drivers (ref include/pipeline/train.hpp:14-136): ``train_semi_async_epoch``, ``train_sync_epoch``,
``validate_semi_async_epoch`` and ``train_model(coordinator, train_loader, test_loader, config)``."""
from __future__ import annotations

import time
from typing import Optional

from ...train import TrainingConfig, split_microbatches
from .coordinator import Coordinator


def train_semi_async_epoch(coordinator: Coordinator, train_loader, config: TrainingConfig):
    train_loader.shuffle()
    train_loader.reset()
    coordinator.set_training(True)
    total_loss, total_correct, samples, batches = 0.0, 0.0, 0, 0
    t_epoch = time.perf_counter()
    while True:
        batch = train_loader.get_next_batch()
        if batch is None:
            break
        data, labels = batch
        m = coordinator.num_microbatches
        xs, ts = split_microbatches(data, m), split_microbatches(labels, m)
        if len(xs) != m:   # batch smaller than the microbatch count: run it as one microbatch
            coordinator.set_num_microbatches(len(xs))
        t0 = time.perf_counter()
        loss = coordinator.async_process_batch(xs, ts)
        res = coordinator.update_parameters()
        coordinator.set_num_microbatches(m)
        if res["loss"] == res["loss"]:
            loss = res["loss"]
        total_loss += loss
        total_correct += res.get("correct", 0.0)
        samples += data.shape[0]
        batches += 1
        if batches % config.progress_print_interval == 0:
            print(f"Batch ID: {batches}, Batch's Loss: {loss:.4f}, Cumulative Accuracy: "
                  f"{100.0 * total_correct / max(1, samples):.2f}%, batch time {1000 * (time.perf_counter() - t0):.1f} ms")
            if config.profiler_type != "NONE":
                coordinator.print_profiling_on_all_stages()
            if config.profiler_type == "NORMAL":
                coordinator.clear_profiling_data()
    return total_loss / max(1, batches), total_correct / max(1, samples), time.perf_counter() - t_epoch


def train_sync_epoch(coordinator: Coordinator, train_loader, config: TrainingConfig):
    """GPipe ordering with the loss on the coordinator (examples/sync_pipeline_coordinator.cpp:99-201).
    Returns (mean loss, accuracy, seconds) like the semi-async epoch."""
    train_loader.shuffle()
    train_loader.reset()
    coordinator.set_training(True)
    keep = coordinator.loss_on_last_stage
    total_loss, total_correct, samples, batches = 0.0, 0.0, 0, 0
    t_epoch = time.perf_counter()
    while True:
        batch = train_loader.get_next_batch()
        if batch is None:
            break
        data, labels = batch
        m = coordinator.num_microbatches
        xs, ts = split_microbatches(data, m), split_microbatches(labels, m)
        coordinator.set_num_microbatches(len(xs))
        total_loss += coordinator.sync_process_batch(xs, ts)
        total_correct += getattr(coordinator, "last_correct", 0.0)
        coordinator.update_parameters()
        coordinator.set_num_microbatches(m)
        samples += data.shape[0]
        batches += 1
    coordinator.loss_on_last_stage = keep
    return total_loss / max(1, batches), total_correct / max(1, samples), time.perf_counter() - t_epoch


def validate_sync_epoch(coordinator: Coordinator, test_loader):
    """Forward-only pass in EVAL_MODE with the loss on the coordinator (the GPipe arrangement of train_sync_epoch)."""
    from .protocol import CommandType
    test_loader.reset()
    coordinator.set_training(False)
    total_loss, total_correct, samples, batches = 0.0, 0.0, 0, 0
    while True:
        batch = test_loader.get_next_batch()
        if batch is None:
            break
        data, labels = batch
        coordinator.forward(data, 0)
        coordinator.join(CommandType.FORWARD_JOB, 1, 120.0)
        for msg in coordinator.comm.dequeue_all_messages_by_type(CommandType.FORWARD_JOB):
            out = msg.payload.data
            loss, _, correct = coordinator.loss_function.forward_backward(out, labels.to(out.device), want_grad=False)
            total_loss += float(loss)
            total_correct += float(correct) if correct is not None else 0.0
        samples += data.shape[0]
        batches += 1
    coordinator.set_training(True)
    return total_loss / max(1, batches), total_correct / max(1, samples)


def validate_semi_async_epoch(coordinator: Coordinator, test_loader):
    """Forward-only pass in EVAL_MODE; the last stage accumulates loss / correct counts, fetched (and reset) with a
    STATUS_REQUEST round-trip -- no optimizer step happens."""
    from .protocol import CommandType
    test_loader.reset()
    coordinator.set_training(False)
    samples = 0
    while True:
        batch = test_loader.get_next_batch()
        if batch is None:
            break
        data, labels = batch
        coordinator.forward(data, 0, labels)
        coordinator.join(CommandType.FORWARD_JOB, 1, 120.0)
        coordinator.comm.dequeue_all_messages_by_type(CommandType.FORWARD_JOB)
        samples += data.shape[0]
    res = coordinator.fetch_loss_stats()
    coordinator.set_training(True)
    return res["loss"], res["correct"] / max(1, samples)


def train_model(coordinator: Coordinator, train_loader, test_loader, config: Optional[TrainingConfig] = None,
                schedule: str = "semi_async"):
    config = config or TrainingConfig()
    coordinator.set_num_microbatches(config.num_microbatches)
    train_loader.prepare_batches(config.batch_size)
    if test_loader is not None:
        test_loader.prepare_batches(config.batch_size)
    history = []
    for epoch in range(config.epochs):
        print(f"Epoch {epoch + 1}/{config.epochs}")
        if schedule == "sync":
            loss, acc, secs = train_sync_epoch(coordinator, train_loader, config)
        else:
            loss, acc, secs = train_semi_async_epoch(coordinator, train_loader, config)
        print(f"Epoch {epoch + 1} - loss {loss:.4f}, accuracy {100 * acc:.2f}%, {secs * 1000:.0f} ms")
        history.append({"epoch": epoch + 1, "loss": loss, "accuracy": acc})
        train_loader.prepare_batches(config.batch_size)
    return history


# ------------------------------------------------------------------------------------------------ recovery
class ResilientBatchRunner:
    """Runs batches through a coordinator and survives the loss of a stage worker.

    Every ``snapshot_every`` completed batches the stages' parameters are gathered into the coordinator's copy of the
    model (``collect_current_parameters``).  When a batch fails because a stage died (connection lost, stage error,
    timeout), ``replacement(stage_index, error)`` is asked for the endpoint of a fresh worker -- start a process, pick a
    spare, whatever the deployment does; it may return None to give up -- the stage is replaced
    (``Coordinator.replace_stage``: every stage reloads the last snapshot) and the batches since that snapshot are run
    again from the replay buffer, so that the sequence of optimizer steps is the one an undisturbed run takes (exactly so
    with plain SGD; the replaced stage's optimizer moments restart from zero otherwise).  The reference has no
    recovery: a lost worker ends the run."""

    def __init__(self, coordinator: Coordinator, replacement, snapshot_every: int = 50, max_recoveries: int = 3,
                 batch_timeout: float = 120.0):
        self.coordinator, self.replacement = coordinator, replacement
        self.snapshot_every, self.max_recoveries = max(1, int(snapshot_every)), int(max_recoveries)
        self.batch_timeout = batch_timeout
        self.recoveries = 0
        self._since_snapshot = []          # (microbatch inputs, microbatch labels) of batches after the last snapshot
        self._have_snapshot = False

    def _lost_stage(self) -> int:
        c = self.coordinator
        alive = getattr(c.comm, "peer_alive", None)
        if alive is not None:
            for i, name in enumerate(c.stage_names):
                if not alive(name):
                    return i
        for i, name in enumerate(c.stage_names):       # a stage that reported its neighbour gone names itself
            if any(f"{name}:" in e and "next_stage" in e for e in c.errors[-4:]):
                return min(i + 1, c.num_stages - 1)
            if any(f"{name}:" in e and "prev_stage" in e for e in c.errors[-4:]):
                return max(i - 1, 0)
        return -1

    def _one(self, xs, ts) -> dict:
        c = self.coordinator
        c.async_process_batch(xs, ts, timeout=self.batch_timeout)
        return c.update_parameters(timeout=self.batch_timeout)

    def snapshot(self) -> bool:
        ok = self.coordinator.collect_current_parameters()
        if ok:
            self._since_snapshot.clear()
            self._have_snapshot = True
        return ok

    def run_batch(self, microbatch_inputs, microbatch_labels) -> dict:
        """One optimisation step; returns ``update_parameters``' result.  Raises if recovery is impossible."""
        if not self._have_snapshot:
            if not self.snapshot():
                raise RuntimeError("cannot take the initial parameter snapshot")
        xs, ts = list(microbatch_inputs), list(microbatch_labels)
        while True:
            try:
                res = self._one(xs, ts)
                break
            except Exception as e:             # PipelineStageLost, ConnectionError, stage ERROR_REPORT, PipelineTimeout
                idx = self._lost_stage()
                if idx < 0 or self.recoveries >= self.max_recoveries:
                    raise
                endpoint = self.replacement(idx, e)
                if endpoint is None:
                    raise
                self.recoveries += 1
                print(f"[pipeline] stage {idx} lost ({type(e).__name__}); replacing it and replaying "
                      f"{len(self._since_snapshot)} batch(es) since the last snapshot")
                if not self.coordinator.replace_stage(idx, endpoint, timeout=self.batch_timeout):
                    raise RuntimeError(f"could not bring a replacement for stage {idx} into the pipeline") from e
                for (px, pt) in self._since_snapshot:
                    self._one(px, pt)
        self._since_snapshot.append((xs, ts))
        if len(self._since_snapshot) >= self.snapshot_every:
            self.snapshot()
        return res


# ------------------------------------------------------------------------------------------------ SPMD engine driver
def train_engine_model(engine, train_loader, test_loader, config: Optional[TrainingConfig] = None, replica: int = 0,
                       num_replicas: int = 1, snapshot: bool = False):
    """Epoch loop for the SPMD :class:`PipelineEngine` (one process per stage x replica, ``torchrun``).

    Every rank iterates the same batches as the other stages of its replica (same seed; replicas take disjoint shards):
    the first stage uses the images, the last stage the labels.  Loss / accuracy live on the last stages; one small
    all-reduce over the whole job per epoch turns them into global means that every rank returns.  Rank 0 prints."""
    import torch
    import torch.distributed as dist
    config = config or TrainingConfig()
    rank = dist.get_rank() if dist.is_initialized() else 0
    say = print if rank == 0 else (lambda *a, **k: None)
    for ld in (train_loader, test_loader):
        ld.set_seed(1234)
        ld.set_shard(replica, num_replicas)
        ld.prepare_batches(config.batch_size)
    dev = engine.device
    red_dev = dev if dev.type == "cuda" else torch.device("cpu")

    def run(loader, step_fn):
        loader.reset()
        acc = torch.zeros(4, dtype=torch.float64, device=red_dev)       # loss sum, correct, samples, batches
        for data, labels in loader:
            loss, correct = step_fn(data, labels)
            if engine.is_last:
                acc[0] += loss.double().to(red_dev)
                if correct is not None:
                    acc[1] += correct.double().to(red_dev)
                acc[2] += data.shape[0]
                acc[3] += 1
        if dist.is_initialized():
            dist.all_reduce(acc)
        return float(acc[0] / acc[3].clamp(min=1)), float(acc[1] / acc[2].clamp(min=1))

    history = []
    best_val_accuracy = 0.0
    for epoch in range(config.epochs):
        say(f"Epoch {epoch + 1}/{config.epochs}")
        t0 = time.perf_counter()
        train_loader.shuffle()
        tl, ta = run(train_loader, engine.train_step)
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        train_ms = (time.perf_counter() - t0) * 1000
        vl, va = run(test_loader, engine.eval_step)
        say("-" * 60)
        say(f"Epoch {epoch + 1}/{config.epochs} completed in {train_ms:.0f}ms")
        say(f"Training   - Loss: {tl:.4f}, Accuracy: {ta * 100:.2f}%")
        say(f"Validation - Loss: {vl:.4f}, Accuracy: {va * 100:.2f}%")
        say("=" * 60)
        history.append({"epoch": epoch + 1, "train_loss": tl, "train_acc": ta, "val_loss": vl, "val_acc": va,
                        "train_ms": train_ms})
        if snapshot and va > best_val_accuracy:
            # every stage takes part in the gather; replica 0's first stage writes <snapshot_dir>/<model name>.json/.bin
            # (+ .state.* with the BatchNorm statistics) -- loadable by Sequential.from_file and by the reference
            full = engine.gather_full_model(0)
            if full is not None and replica == 0:
                import os
                path = os.path.join(config.snapshot_dir, engine.full_model.name())
                full.save_to_file(path, save_state=True)
                say(f"New best validation accuracy: {va * 100:.2f}% -- model saved to {path}")
        best_val_accuracy = max(best_val_accuracy, va)
        if (epoch + 1) % config.lr_decay_interval == 0:
            cur = engine.optimizer.get_learning_rate()
            engine.optimizer.set_learning_rate(cur * config.lr_decay_factor)
        train_loader.prepare_batches(config.batch_size)      # re-apply augmentation, like the single-device loop
    return history
