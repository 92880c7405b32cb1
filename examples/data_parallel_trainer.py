#!/usr/bin/env python
"""Data-parallel trainer (new capability: the reference has no data parallelism).  One process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29500 \\
        examples/data_parallel_trainer.py --model resnet50_tiny_imagenet --device-type GPU --batch-size 256

Every replica trains on its shard of the dataset (BATCH_SIZE is the per-replica batch); gradients are reduced, the
optimizer applied and the parameters re-broadcast by ONE kernel over NVLink peer memory, BatchNorm statistics are merged
across replicas inside the BN kernels.  With DEVICE_TYPE=CPU the same script runs on gloo (plumbing / CI).
MODEL: any key of dcnn_b200.models.MODEL_ZOO; the dataset follows the model name (mnist / cifar10 / cifar100 /
tiny_imagenet), synthetic data of that shape when the dataset is not on disk."""
import os

import torch
import torch.distributed as dist
from _common import get_env, loaders, setup

from dcnn_b200 import nn
from dcnn_b200.data import (create_cifar10_data_loaders, create_cifar100_data_loaders, create_mnist_data_loaders_dir,
                            create_tiny_imagenet_data_loaders)
from dcnn_b200.models import MODEL_ZOO
from dcnn_b200.parallel.dp import DataParallel
from dcnn_b200.train import train_classification_model

DATASETS = {
    "mnist": (create_mnist_data_loaders_dir, "./data/mnist", (1, 28, 28), 10),
    "cifar100": (create_cifar100_data_loaders, "./data/cifar-100-binary", (3, 32, 32), 100),
    "cifar10": (create_cifar10_data_loaders, "./data/cifar-10-batches-bin", (3, 32, 32), 10),
    "tiny_imagenet": (create_tiny_imagenet_data_loaders, "./data/tiny-imagenet-200", (3, 64, 64), 200),
}


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if rank != 0:                                   # only replica 0 narrates the set-up
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            cfg = setup()
    else:
        cfg = setup()
    name = get_env("MODEL", "resnet9_cifar10")
    if name not in MODEL_ZOO:
        raise SystemExit(f"unknown MODEL '{name}'; choose from {sorted(MODEL_ZOO)}")
    key = next((k for k in ("cifar100", "cifar10", "tiny_imagenet", "mnist") if k in name), None)
    if key is None:
        raise SystemExit(f"no dataset is associated with model '{name}' (ImageNet-scale loaders are not part of the port)")
    factory, data_dir, shape, classes = DATASETS[key]
    gpu = cfg.device_type == "GPU"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if gpu:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    train_loader, test_loader = loaders(factory, data_dir, shape, classes, cfg,
                                        n_train=get_env("SYNTHETIC_TRAIN_SAMPLES", cfg.batch_size * 8 * world),
                                        n_test=get_env("SYNTHETIC_TEST_SAMPLES", cfg.batch_size * 2 * world))
    model = MODEL_ZOO[name]()
    model.set_seed(42)                              # identical initial weights on every replica
    model.set_device(torch.device("cuda", local) if gpu else torch.device("cpu"))
    model.initialize()
    optimizer = nn.Adam(get_env("LR_INITIAL", 0.001), 0.9, 0.999, 1e-8, get_env("WEIGHT_DECAY", 0.0))
    dp = DataParallel(model, optimizer, dist.group.WORLD, sync_bn=get_env("SYNC_BN", True))
    if rank == 0:
        print(f"Data parallel: {world} replica(s), per-replica batch {cfg.batch_size}, global batch {cfg.batch_size * world}")
    history = train_classification_model(model, train_loader, test_loader, optimizer,
                                         nn.LossFactory.create_logsoftmax_crossentropy(), cfg, data_parallel=dp)
    if rank == 0 and history:
        print(f"Final: train loss {history[-1]['train_loss']:.4f}, validation accuracy {history[-1]['val_acc'] * 100:.2f}%")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
