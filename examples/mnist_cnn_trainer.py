#!/usr/bin/env python
"""MNIST CNN trainer -- counterpart of the reference's examples/mnist_cnn_trainer.cpp (same model, optimizer, loss and augmentation;
configuration from ./.env / environment: EPOCHS, BATCH_SIZE, LR_INITIAL, DEVICE_TYPE, ...)."""
from _common import get_env, loaders, setup

from dcnn_b200 import nn
from dcnn_b200.data import AugmentationBuilder, create_mnist_data_loaders_dir
from dcnn_b200.models import create_mnist_trainer
from dcnn_b200.train import train_classification_model


def main():
    cfg = setup()
    lr_initial = get_env("LR_INITIAL", 0.01)
    train_loader, test_loader = loaders(create_mnist_data_loaders_dir, "./data/mnist", (1, 28, 28), 10, cfg)
    print(f"Successfully loaded training data: {train_loader.size()} samples")
    print(f"Successfully loaded test data: {test_loader.size()} samples")
    train_loader.set_augmentation(AugmentationBuilder().contrast(0.3, 0.15).gaussian_noise(0.3, 0.05).build())
    val_aug = None
    if val_aug is not None:
        test_loader.set_augmentation(val_aug)
    model = create_mnist_trainer()
    model.set_device(cfg.torch_device())
    model.initialize()
    optimizer = nn.Adam(lr_initial, 0.9, 0.999, 1e-8)
    loss_function = nn.LossFactory.create_logsoftmax_crossentropy()
    train_classification_model(model, train_loader, test_loader, optimizer, loss_function, cfg)


if __name__ == "__main__":
    main()
