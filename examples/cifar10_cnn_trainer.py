#!/usr/bin/env python
"""CIFAR-10 CNN trainer -- counterpart of the reference's examples/cifar10_cnn_trainer.cpp (same model, optimizer, loss and augmentation;
configuration from ./.env / environment: EPOCHS, BATCH_SIZE, LR_INITIAL, DEVICE_TYPE, ...)."""
from _common import get_env, loaders, setup

from dcnn_b200 import nn
from dcnn_b200.data import AugmentationBuilder, create_cifar10_data_loaders
from dcnn_b200.models import create_cifar10_trainer_v2
from dcnn_b200.train import train_classification_model


def main():
    cfg = setup()
    lr_initial = get_env("LR_INITIAL", 0.001)
    train_loader, test_loader = loaders(create_cifar10_data_loaders, "./data/cifar-10-batches-bin", (3, 32, 32), 10, cfg)
    print(f"Successfully loaded training data: {train_loader.size()} samples")
    print(f"Successfully loaded test data: {test_loader.size()} samples")
    train_loader.set_augmentation(AugmentationBuilder().horizontal_flip(0.5).random_crop(0.5, 4).normalize((0.49139968, 0.48215827, 0.44653124), (0.24703233, 0.24348505, 0.26158768)).build())
    val_aug = AugmentationBuilder().normalize((0.49139968, 0.48215827, 0.44653124), (0.24703233, 0.24348505, 0.26158768)).build()
    if val_aug is not None:
        test_loader.set_augmentation(val_aug)
    model = create_cifar10_trainer_v2()
    model.set_device(cfg.torch_device())
    model.initialize()
    optimizer = nn.Adam(lr_initial, 0.9, 0.999, 1e-8)
    loss_function = nn.LossFactory.create_logsoftmax_crossentropy()
    train_classification_model(model, train_loader, test_loader, optimizer, loss_function, cfg)


if __name__ == "__main__":
    main()
