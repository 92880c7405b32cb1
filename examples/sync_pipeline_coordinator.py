#!/usr/bin/env python
"""Synchronous (GPipe) pipeline coordinator, MNIST CNN, 2 stages -- counterpart of the reference's examples/sync_pipeline_coordinator.cpp.
Single process: stages are threads (one per GPU when DEVICE_TYPE=GPU).  Multi-process:
torchrun --nproc-per-node <S+1> examples/sync_pipeline_coordinator.py (rank 0 = coordinator, ranks 1..S = stage workers)."""
from _common import get_env, loaders, setup
from _pipeline_common import make_coordinator

from dcnn_b200 import nn
from dcnn_b200.data import create_mnist_data_loaders_dir
from dcnn_b200.models import create_mnist_trainer
from dcnn_b200.parallel.pipeline.train import train_model, validate_sync_epoch
from dcnn_b200.partitioner import NaivePartitioner


def main():
    cfg = setup()
    model = create_mnist_trainer()                      # NOT initialised on the coordinator: stages build their own shards
    optimizer = nn.Adam(get_env('LR_INITIAL', 0.01))
    loss = nn.LossFactory.create_softmax_crossentropy()   # the reference names CrossEntropy; the model emits logits
    # GPipe ordering computes the loss on the coordinator, like the reference example
    coord, is_coord = make_coordinator(model, optimizer, loss, get_env("NUM_STAGES", 2), cfg, NaivePartitioner())
    if not is_coord:
        return
    coord.loss_on_last_stage = False
    train_loader, test_loader = loaders(create_mnist_data_loaders_dir, "./data/mnist", (1, 28, 28), 10, cfg)
    history = train_model(coord, train_loader, test_loader, cfg, schedule="sync")
    val_loss, val_acc = validate_sync_epoch(coord, test_loader)
    print(f"Validation - Loss: {val_loss:.4f}, Accuracy: {100 * val_acc:.2f}%")
    coord.print_profiling_on_all_stages()
    if coord.save_checkpoint("model_snapshots/" + model.name()):
        print("Model saved to model_snapshots/" + model.name())
    coord.stop()
    return history


if __name__ == "__main__":
    main()
