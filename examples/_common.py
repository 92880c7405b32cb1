"""Shared scaffolding of the example entrypoints (the reference's examples/*.cpp): .env + environment configuration,
device selection, dataset-or-synthetic loaders.  ``SYNTHETIC=1`` (or a missing dataset directory) substitutes synthetic
data of the dataset's shape, which is what the benchmark configs in BASELINE.json use."""
from __future__ import annotations

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from dcnn_b200.data import SyntheticDataLoader  # noqa: E402
from dcnn_b200.train import TrainingConfig  # noqa: E402
from dcnn_b200.utils.env import get_env, load_env_file, override  # noqa: E402


def parse_flags(argv=None):
    """``--epochs 3 --batch-size=64 --device-type GPU --synthetic`` -> ({"EPOCHS": "3", "BATCH_SIZE": "64",
    "DEVICE_TYPE": "GPU", "SYNTHETIC": "1"}, [positional arguments])."""
    argv = list(sys.argv[1:] if argv is None else argv)
    flags, rest, i = {}, [], 0
    while i < len(argv):
        a = argv[i]
        if a.startswith("--") and len(a) > 2:
            key, eq, val = a[2:].partition("=")
            if not eq:
                if i + 1 < len(argv) and not argv[i + 1].startswith("--"):
                    val, i = argv[i + 1], i + 1
                else:
                    val = "1"            # bare switch
            flags[key.upper().replace("-", "_")] = val
        else:
            rest.append(a)
        i += 1
    return flags, rest


def apply_flag_overrides(argv=None):
    """A flag layer over the reference's .env keys: flags beat the .env file and the inherited
    environment.  Returns the positional arguments that were not consumed."""
    flags, rest = parse_flags(argv)
    for k, v in flags.items():
        override(k, v)
    return rest


def setup():
    print("Loading environment variables...")
    if not (os.path.exists("./.env") and load_env_file("./.env")):
        print("No .env file found, using default training parameters.")
    apply_flag_overrides()               # after the .env file so that command-line flags win
    cfg = TrainingConfig().load_from_env()
    if cfg.device_type == "GPU" and not torch.cuda.is_available():
        raise SystemExit("DEVICE_TYPE=GPU but no CUDA device is visible")
    print(f"Using device type: {cfg.device_type}")
    cfg.print_config()
    return cfg


def loaders(real_factory, data_dir, shape, classes, cfg, n_train=None, n_test=None):
    """Real dataset if present (and SYNTHETIC != 1), else synthetic tensors of the same shape."""
    synthetic = get_env("SYNTHETIC", False) or not os.path.exists(data_dir)
    if not synthetic:
        try:
            return real_factory(data_dir)
        except Exception as e:  # missing/corrupt files: fall back, loudly
            print(f"Could not load dataset from {data_dir} ({e}); using synthetic data")
    n_train = n_train or get_env("SYNTHETIC_TRAIN_SAMPLES", cfg.batch_size * 8)
    n_test = n_test or get_env("SYNTHETIC_TEST_SAMPLES", cfg.batch_size * 2)
    print(f"Using synthetic data: {n_train} train / {n_test} test samples of shape {tuple(shape)}, {classes} classes")
    return SyntheticDataLoader(n_train, shape, classes, seed=0), SyntheticDataLoader(n_test, shape, classes, seed=1)
