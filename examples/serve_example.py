#!/usr/bin/env python
"""Serving a trained snapshot: load <name>.json/.bin, fold the BatchNorms into the convolutions, and answer requests of
any size through a fixed-batch predictor (one CUDA-graph replay per batch on a GPU).  No counterpart in the reference,
whose only inference entry point is the evaluation loop of mnist_cnn_test.

    python examples/mnist_cnn_trainer.py            # writes model_snapshots/mnist_cnn_model.{json,bin,state.*}
    python examples/serve_example.py [snapshot path] [--batch 64] [--requests 20]"""
import argparse
import os
import time

from _common import sys  # noqa: F401  (path setup)

import torch

from dcnn_b200 import nn
from dcnn_b200.models import create_mnist_trainer
from dcnn_b200.serve import Predictor


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("snapshot", nargs="?", default="model_snapshots/mnist_cnn_model")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--requests", type=int, default=20)
    args = ap.parse_args()
    device = "GPU" if torch.cuda.is_available() and os.environ.get("DEVICE_TYPE", "GPU") == "GPU" else "CPU"
    if os.path.exists(args.snapshot + ".json"):
        model = nn.Sequential.from_file(args.snapshot, device=device)
        print(f"Loaded {args.snapshot}")
    else:
        print(f"No snapshot at {args.snapshot}; serving a freshly initialised MNIST CNN")
        model = create_mnist_trainer()
        model.set_device(device)
        model.initialize()
    shape = (1, 28, 28)
    predictor = Predictor(model, shape, batch_size=args.batch)
    g = torch.Generator().manual_seed(0)
    sizes = [int(torch.randint(1, 3 * args.batch, (1,), generator=g)) for _ in range(args.requests)]
    predictor.predict(torch.rand(args.batch, *shape))                      # warm-up (and graph capture on a GPU)
    t0 = time.perf_counter()
    total = 0
    for n in sizes:
        classes = predictor.predict_classes(torch.rand(n, *shape, generator=g))
        total += int(classes.numel())
    if torch.cuda.is_available() and device == "GPU":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(predictor.summary())
    print(f"Served {args.requests} requests / {total} images in {dt * 1e3:.1f} ms ({total / dt:,.0f} images/s)")


if __name__ == "__main__":
    main()
