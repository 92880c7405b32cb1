#!/usr/bin/env python
"""Tiny-ImageNet ResNet-34 trainer -- counterpart of the reference's examples/tiny_imagenet_resnet34.cpp (same model, optimizer, loss and augmentation;
configuration from ./.env / environment: EPOCHS, BATCH_SIZE, LR_INITIAL, DEVICE_TYPE, ...)."""
from _common import get_env, loaders, setup

from dcnn_b200 import nn
from dcnn_b200.data import AugmentationBuilder, create_tiny_imagenet_data_loaders
from dcnn_b200.models import create_resnet34_tiny_imagenet
from dcnn_b200.train import train_classification_model


def main():
    cfg = setup()
    lr_initial = get_env("LR_INITIAL", 0.0001)
    train_loader, test_loader = loaders(create_tiny_imagenet_data_loaders, "./data/tiny-imagenet-200", (3, 64, 64), 200, cfg)
    print(f"Successfully loaded training data: {train_loader.size()} samples")
    print(f"Successfully loaded test data: {test_loader.size()} samples")
    train_loader.set_augmentation(AugmentationBuilder().random_crop(1.0, 4).rotation(0.25, 5.0).horizontal_flip(0.5).brightness(0.2).normalize((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)).build())
    val_aug = AugmentationBuilder().normalize((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)).build()
    if val_aug is not None:
        test_loader.set_augmentation(val_aug)
    model = create_resnet34_tiny_imagenet()
    model.set_device(cfg.torch_device())
    model.initialize()
    optimizer = nn.Adam(lr_initial, 0.9, 0.999, 1e-7)
    loss_function = nn.LossFactory.create_logsoftmax_crossentropy()
    train_classification_model(model, train_loader, test_loader, optimizer, loss_function, cfg)


if __name__ == "__main__":
    main()
