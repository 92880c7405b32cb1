#!/usr/bin/env python
"""Plain-PyTorch trainer for the same models and data -- counterpart of the reference's examples/torch_mnist_trainer.py,
torch_cifar10_trainer.py and torch_tiny_imagenet_trainer.py (its PyTorch comparison scripts).  The network is the zoo
model exported with ``dcnn_b200.io.to_torch_module`` (identical architecture and initial weights), trained with
autograd + ``torch.optim.Adam``; data comes from this library's loaders (or synthetic tensors).

    python examples/torch_trainer.py --model mnist_cnn            # | resnet9_cifar10 | resnet18_tiny_imagenet | ...
Environment / flags as in the other examples (EPOCHS, BATCH_SIZE, LR_INITIAL, DEVICE_TYPE, SYNTHETIC)."""
import time

from _common import get_env, loaders, setup

import torch

from dcnn_b200.data import (create_cifar10_data_loaders, create_cifar100_data_loaders, create_mnist_data_loaders_dir,
                            create_tiny_imagenet_data_loaders)
from dcnn_b200.io import to_torch_module
from dcnn_b200.models import MODEL_ZOO

DATASETS = {"mnist": (create_mnist_data_loaders_dir, "./data/mnist", (1, 28, 28), 10),
            "cifar100": (create_cifar100_data_loaders, "./data/cifar-100-binary", (3, 32, 32), 100),
            "cifar10": (create_cifar10_data_loaders, "./data/cifar-10-batches-bin", (3, 32, 32), 10),
            "tiny_imagenet": (create_tiny_imagenet_data_loaders, "./data/tiny-imagenet-200", (3, 64, 64), 200)}


def main():
    cfg = setup()
    name = get_env("MODEL", "mnist_cnn")
    if name not in MODEL_ZOO:
        raise SystemExit(f"--model {name}: not in MODEL_ZOO ({sorted(MODEL_ZOO)})")
    key = next((k for k in DATASETS if k in name), None)
    if key is None:
        raise SystemExit(f"no dataset is associated with model {name}")
    factory, path, shape, classes = DATASETS[key]
    train_loader, test_loader = loaders(factory, path, shape, classes, cfg)
    dev = cfg.torch_device()
    model = MODEL_ZOO[name]()
    model.set_seed(1)
    model.initialize()
    net = to_torch_module(model).to(dev)
    if dev.type == "cuda":
        net = net.to(memory_format=torch.channels_last)
        torch.backends.cudnn.benchmark = True
    opt = torch.optim.Adam(net.parameters(), lr=get_env("LR_INITIAL", 0.001))
    train_loader.prepare_batches(cfg.batch_size)
    test_loader.prepare_batches(cfg.batch_size)
    for epoch in range(cfg.epochs):
        net.train()
        train_loader.shuffle()
        train_loader.reset()
        t0, seen, correct, loss_sum, batches = time.perf_counter(), 0, 0, 0.0, 0
        while (batch := train_loader.get_next_batch()) is not None:
            x, t = batch[0].to(dev, non_blocking=True), batch[1].to(dev, non_blocking=True)
            labels = t.reshape(t.shape[0], -1).argmax(1)
            opt.zero_grad(set_to_none=True)
            logits = net(x).flatten(1)
            loss = torch.nn.functional.cross_entropy(logits, labels)
            loss.backward()
            opt.step()
            loss_sum += float(loss.detach())
            correct += int((logits.argmax(1) == labels).sum())
            seen += x.shape[0]
            batches += 1
        dt = time.perf_counter() - t0
        net.eval()
        test_loader.reset()
        v_loss, v_correct, v_seen, v_batches = 0.0, 0, 0, 0
        with torch.no_grad():
            while (batch := test_loader.get_next_batch()) is not None:
                x, t = batch[0].to(dev), batch[1].to(dev)
                labels = t.reshape(t.shape[0], -1).argmax(1)
                logits = net(x).flatten(1)
                v_loss += float(torch.nn.functional.cross_entropy(logits, labels))
                v_correct += int((logits.argmax(1) == labels).sum())
                v_seen += x.shape[0]
                v_batches += 1
        print(f"Epoch {epoch + 1}/{cfg.epochs} completed in {dt * 1e3:.0f}ms ({seen / dt:,.0f} images/s)")
        print(f"Training   - Loss: {loss_sum / max(1, batches):.4f}, Accuracy: {100.0 * correct / max(1, seen):.2f}%")
        print(f"Validation - Loss: {v_loss / max(1, v_batches):.4f}, Accuracy: {100.0 * v_correct / max(1, v_seen):.2f}%")


if __name__ == "__main__":
    main()
