#!/usr/bin/env python
"""Load a saved MNIST snapshot and evaluate it -- counterpart of examples/mnist_cnn_test.cpp."""
import sys

from _common import loaders, setup

from dcnn_b200 import Sequential, nn
from dcnn_b200.data import create_mnist_data_loaders_dir
from dcnn_b200.train import validate_class_model


def main():
    cfg = setup()
    path = sys.argv[1] if len(sys.argv) > 1 else "model_snapshots/mnist_cnn_model"
    model = Sequential.from_file(path, device=cfg.torch_device())
    _, test_loader = loaders(create_mnist_data_loaders_dir, "./data/mnist", (1, 28, 28), 10, cfg)
    test_loader.prepare_batches(cfg.batch_size)
    loss, acc = validate_class_model(model, test_loader, nn.LossFactory.create_logsoftmax_crossentropy())
    print(f"Test - Loss: {loss:.4f}, Accuracy: {acc * 100:.2f}%")


if __name__ == "__main__":
    main()
