from .augmentation import *  # noqa: F401,F403
from .augmentation import AugmentationBuilder, AugmentationStrategy  # noqa: F401
from .loaders import (BaseDataLoader, CIFAR10DataLoader, CIFAR100DataLoader, ImageDataLoader, MNISTDataLoader,  # noqa: F401
                      RegressionDataLoader, SyntheticDataLoader, TinyImageNetDataLoader, WiFiDataLoader,
                      create_cifar10_data_loaders, create_cifar100_data_loaders, create_mnist_data_loaders, create_mnist_data_loaders_dir,
                      create_tiny_imagenet_data_loaders, create_data_loader, normalize_to_unit_range, train_val_split)
from .prefetch import PrefetchLoader  # noqa: F401
