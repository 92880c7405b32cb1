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
    """GPipe ordering with the loss on the coordinator (examples/sync_pipeline_coordinator.cpp:99-201)."""
    train_loader.shuffle()
    train_loader.reset()
    coordinator.set_training(True)
    keep = coordinator.loss_on_last_stage
    total_loss, batches = 0.0, 0
    while True:
        batch = train_loader.get_next_batch()
        if batch is None:
            break
        data, labels = batch
        m = coordinator.num_microbatches
        xs, ts = split_microbatches(data, m), split_microbatches(labels, m)
        coordinator.set_num_microbatches(len(xs))
        total_loss += coordinator.sync_process_batch(xs, ts)
        coordinator.update_parameters()
        coordinator.set_num_microbatches(m)
        batches += 1
    coordinator.loss_on_last_stage = keep
    return total_loss / max(1, batches)


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
            loss = train_sync_epoch(coordinator, train_loader, config)
            acc, secs = float("nan"), 0.0
        else:
            loss, acc, secs = train_semi_async_epoch(coordinator, train_loader, config)
        print(f"Epoch {epoch + 1} - loss {loss:.4f}, accuracy {100 * acc:.2f}%, {secs * 1000:.0f} ms")
        history.append({"epoch": epoch + 1, "loss": loss, "accuracy": acc})
        train_loader.prepare_batches(config.batch_size)
    return history
