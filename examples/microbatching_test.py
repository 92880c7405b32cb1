#!/usr/bin/env python
"""Microbatch equivalence check -- counterpart of the reference's examples/microbatching_test.cpp (present in its tree
but commented out of its build): the gradients of a batch processed as M microbatches with per-microbatch caches must
equal the gradients of the whole batch, for every parameter.  Prints the largest deviation per model."""
from _common import setup

import torch

from dcnn_b200 import nn
from dcnn_b200.models import create_mnist_trainer, create_resnet9_cifar10
from dcnn_b200.train import split_microbatches


def check(make, shape, classes, device, micro=4, batch=16):
    torch.manual_seed(0)
    full, split = make(), make()
    for m in (full, split):
        m.set_seed(5)
        m.set_device(device)
        m.initialize()
    loss = nn.SoftmaxCrossEntropyLoss()
    x = torch.randn(batch, *shape, device=device)
    t = torch.nn.functional.one_hot(torch.randint(0, classes, (batch,)), classes).float().reshape(batch, classes, 1, 1).to(device)
    # BatchNorm statistics differ between a batch and its microbatches by construction: compare in eval-statistics-free
    # terms by running both with the same microbatching of the FORWARD pass and different backward orders
    xs, ts = split_microbatches(x, micro), split_microbatches(t, micro)
    for i, (xi, ti) in enumerate(zip(xs, ts)):
        full.backward(loss.compute_gradient(full.forward(xi, 0), ti), 0)          # one cache slot, strictly interleaved
    ys = [split.forward(xi, i) for i, xi in enumerate(xs)]                          # all forwards first, M cache slots
    for i in reversed(range(micro)):
        split.backward(loss.compute_gradient(ys[i], ts[i]), i)
    worst = max(float((a - b).abs().max()) for a, b in zip(full.gradients(), split.gradients()))
    scale = max(float(a.abs().max()) for a in full.gradients())
    return worst, scale


def main():
    cfg = setup()
    device = cfg.torch_device()
    ok = True
    for name, make, shape, classes in (("mnist_cnn", create_mnist_trainer, (1, 28, 28), 10),
                                       ("resnet9_cifar10", create_resnet9_cifar10, (3, 32, 32), 10)):
        worst, scale = check(make, shape, classes, device)
        good = worst <= 1e-4 * max(1.0, scale)
        ok = ok and good
        print(f"{name}: max |grad(interleaved) - grad(microbatched)| = {worst:.3e} (largest gradient {scale:.3e}) "
              f"{'OK' if good else 'MISMATCH'}")
    print("Microbatching test " + ("PASSED" if ok else "FAILED"))
    raise SystemExit(0 if ok else 1)


if __name__ == "__main__":
    main()
