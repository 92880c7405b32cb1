"""``python -m dcnn_b200 <command>`` -- a small front door to the package.

    info                       version, library paths, devices, CPU features
    build [--force]            compile libdcnn_b200.so (nvcc, sm_100a) and libdcnn_host.so (g++) in-tree
    models                     the model zoo with parameter counts
    summary --model NAME [--batch N]     per-layer shapes and FLOPs of a zoo model
    train --model NAME [--epochs E --batch-size B --lr LR --device CPU|GPU --data DIR --synthetic --snapshot PATH]
                               train a zoo model on its dataset (synthetic data of the right shape without one)
    predict --snapshot PATH --input N,C,H,W [--device CPU|GPU]     load a snapshot and time the folded predictor

The reference ships one executable per model; the examples/ directory keeps those.  This command covers the common
cases without writing a script."""
from __future__ import annotations

import argparse
import os
import sys
import time


def _info(_):
    import torch
    import dcnn_b200
    from dcnn_b200 import _hostlib, _lib
    from dcnn_b200.utils.device import DeviceManager
    from dcnn_b200.utils.hardware import HardwareInfo
    print(f"dcnn_b200 {dcnn_b200.__version__}  (torch {torch.__version__}, CUDA available: {torch.cuda.is_available()})")
    for name, path in (("device library", _lib.LIB_PATH), ("host library", _hostlib.PATH)):
        print(f"  {name}: {path} [{'built' if os.path.exists(path) else 'not built: python -m dcnn_b200 build'}]")
    print("  devices:", ", ".join(DeviceManager.get_instance().get_available_device_ids()))
    hw = HardwareInfo()
    print(f"  cpu: {hw.get_model_name()} | {hw.get_physical_cores()} cores / {hw.get_logical_cores()} threads | "
          f"AVX2 {hw.supports_avx2()} AVX-512 {hw.supports_avx512()} | "
          f"memory {hw.get_total_memory_bytes() / 2**30:.1f} GiB")
    for g in hw.info.get("gpus", []):
        print(f"  gpu {g['index']}: {g['name']}, {g['sms']} SMs, {g['memory_bytes'] / 2**30:.0f} GiB, sm_{g['cc']}")
    return 0


def _build(args):
    from dcnn_b200.build import build, build_host
    print(build_host(force=args.force))
    print(build(force=args.force))
    return 0


def _zoo(name):
    from dcnn_b200.models import MODEL_ZOO
    if name not in MODEL_ZOO:
        raise SystemExit(f"unknown model '{name}'; choose from: {', '.join(sorted(MODEL_ZOO))}")
    return MODEL_ZOO[name]()


def _dataset_for(name):
    from dcnn_b200.data import (create_cifar10_data_loaders, create_cifar100_data_loaders,
                                create_mnist_data_loaders_dir, create_tiny_imagenet_data_loaders)
    table = {"mnist": (create_mnist_data_loaders_dir, "./data/mnist", (1, 28, 28), 10),
             "cifar100": (create_cifar100_data_loaders, "./data/cifar-100-binary", (3, 32, 32), 100),
             "cifar10": (create_cifar10_data_loaders, "./data/cifar-10-batches-bin", (3, 32, 32), 10),
             "tiny_imagenet": (create_tiny_imagenet_data_loaders, "./data/tiny-imagenet-200", (3, 64, 64), 200),
             "imagenet": (None, "", (3, 224, 224), 1000)}
    for key in ("cifar100", "cifar10", "tiny_imagenet", "imagenet", "mnist"):
        if key in name:
            return table[key]
    raise SystemExit(f"no dataset shape is associated with model '{name}'")


def _models(_):
    from dcnn_b200.models import MODEL_ZOO
    for name in sorted(MODEL_ZOO):
        m = MODEL_ZOO[name]()
        m.initialize()
        n = sum(p.numel() for p in m.parameters())
        print(f"{name:28s} {m.layer_size():3d} top-level layers  {n / 1e6:8.2f} M parameters")
    return 0


def _summary(args):
    m = _zoo(args.model)
    _, _, shape, _ = _dataset_for(args.model)
    m.print_summary([args.batch, *shape])
    return 0


def _train(args):
    import torch
    from dcnn_b200 import nn
    from dcnn_b200.data import SyntheticDataLoader
    from dcnn_b200.train import TrainingConfig, train_classification_model
    factory, default_dir, shape, classes = _dataset_for(args.model)
    cfg = TrainingConfig().load_from_env()
    cfg.epochs, cfg.batch_size, cfg.device_type = args.epochs, args.batch_size, args.device.upper()
    if cfg.device_type == "GPU" and not torch.cuda.is_available():
        raise SystemExit("--device GPU but no CUDA device is visible")
    data_dir = args.data or default_dir
    if args.synthetic or factory is None or not os.path.exists(data_dir):
        n = args.batch_size * 8
        print(f"Using synthetic data: {n} train / {args.batch_size * 2} test samples of shape {shape}")
        train, test = SyntheticDataLoader(n, shape, classes, 0), SyntheticDataLoader(args.batch_size * 2, shape, classes, 1)
    else:
        train, test = factory(data_dir)
    model = _zoo(args.model)
    model.set_device(cfg.torch_device())
    model.initialize()
    opt = nn.Adam(args.lr)
    hist = train_classification_model(model, train, test, opt, nn.LossFactory.create_logsoftmax_crossentropy(), cfg)
    if args.snapshot:
        model.save_to_file(args.snapshot, save_state=True)
        print(f"Snapshot written to {args.snapshot}.json / .bin")
    return 0 if hist is not None else 1


def _predict(args):
    import torch
    from dcnn_b200 import nn
    from dcnn_b200.serve import Predictor
    n, c, h, w = (int(v) for v in args.input.split(","))
    dev = "GPU" if args.device.upper() == "GPU" and torch.cuda.is_available() else "CPU"
    model = nn.Sequential.from_file(args.snapshot, device=dev)
    p = Predictor(model, (c, h, w), batch_size=n)
    x = torch.rand(n, c, h, w)
    p.predict(x)
    t0 = time.perf_counter()
    for _ in range(args.repeat):
        y = p.predict(x)
    if dev == "GPU":
        torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.repeat
    print(p.summary())
    print(f"output {tuple(y.shape)}; {dt * 1e3:.2f} ms per batch of {n} ({n / dt:,.0f} images/s, host-to-host)")
    return 0


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m dcnn_b200", description=__doc__.split("\n\n")[0])
    sub = ap.add_subparsers(dest="cmd", required=True)
    sub.add_parser("info").set_defaults(fn=_info)
    b = sub.add_parser("build")
    b.add_argument("--force", action="store_true")
    b.set_defaults(fn=_build)
    sub.add_parser("models").set_defaults(fn=_models)
    s = sub.add_parser("summary")
    s.add_argument("--model", required=True)
    s.add_argument("--batch", type=int, default=1)
    s.set_defaults(fn=_summary)
    t = sub.add_parser("train")
    t.add_argument("--model", required=True)
    t.add_argument("--epochs", type=int, default=1)
    t.add_argument("--batch-size", type=int, default=64)
    t.add_argument("--lr", type=float, default=1e-3)
    t.add_argument("--device", default="CPU")
    t.add_argument("--data", default="")
    t.add_argument("--synthetic", action="store_true")
    t.add_argument("--snapshot", default="")
    t.set_defaults(fn=_train)
    p = sub.add_parser("predict")
    p.add_argument("--snapshot", required=True)
    p.add_argument("--input", required=True, help="N,C,H,W")
    p.add_argument("--device", default="CPU")
    p.add_argument("--repeat", type=int, default=10)
    p.set_defaults(fn=_predict)
    args = ap.parse_args(argv)
    return args.fn(args)


if __name__ == "__main__":
    sys.exit(main())
