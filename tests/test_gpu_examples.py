"""The product entrypoints on a GPU (counterparts of the reference's examples/*.cpp binaries run with DEVICE_TYPE=GPU,
examples/tiny_imagenet_resnet50.cpp:85-100 + include/nn/train.hpp:202-300): loader -> augmentation -> pinned H2D ->
training step (CUDA graph) -> validation -> snapshot, on synthetic data of the data set's shape.  The snapshot each one
writes is then loaded by the UNMODIFIED reference (baseline/_ref/bin/ref_snapshot read) and pushed through its forward
pass, when the reference subset build is present.  Pipeline coordinators run on >= 2 GPUs over NCCL."""
import os
import struct
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "bin", "ref_snapshot")
ENV = dict(os.environ, EPOCHS="1", BATCH_SIZE="64", SYNTHETIC="1", SYNTHETIC_TRAIN_SAMPLES="256",
           SYNTHETIC_TEST_SAMPLES="2048", DEVICE_TYPE="GPU", NUM_THREADS="4", PROGRESS_PRINT_INTERVAL="1000",
           NUM_MICROBATCHES="2", STEP_TIMEOUT_S="120")

# script, model name of its snapshot, input shape; the ResNets that the reference harness can rebuild are checked with it
TRAINERS = [
    ("mnist_cnn_trainer.py", "mnist_cnn_model", (4, 1, 28, 28)),
    ("cifar10_resnet9.py", None, (4, 3, 32, 32)),
    ("tiny_imagenet_resnet18.py", None, (4, 3, 64, 64)),
    ("tiny_imagenet_resnet34.py", None, (4, 3, 64, 64)),
    ("tiny_imagenet_resnet50.py", None, (4, 3, 64, 64)),
]


@pytest.mark.parametrize("script,snap,shape", TRAINERS, ids=[t[0] for t in TRAINERS])
def test_trainer_runs_on_the_gpu_and_its_snapshot_loads_in_the_reference(tmp_path, script, snap, shape):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", script)], cwd=str(tmp_path), env=ENV,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "Validation - Loss" in r.stdout and "Using device type: GPU" in r.stdout, r.stdout[-2000:]
    snaps = [f[:-5] for f in os.listdir(tmp_path / "model_snapshots") if f.endswith(".json") and ".state" not in f] \
        if (tmp_path / "model_snapshots").exists() else []
    assert snaps, "the trainer wrote no snapshot: " + r.stdout[-1500:]
    prefix = str(tmp_path / "model_snapshots" / (snap or snaps[0]))
    assert os.path.getsize(prefix + ".bin") > 0
    if not os.path.exists(REF):
        pytest.skip("reference subset build absent: snapshot not cross-loaded")
    # the reference loads our snapshot and runs its own training-mode forward; ours (CPU FP32) must agree
    from dcnn_b200.nn import Sequential
    torch.manual_seed(0)
    x = torch.rand(*shape)
    io_path, out_path = str(tmp_path / "in.io.bin"), str(tmp_path / "out.bin")
    with open(io_path, "wb") as f:
        f.write(struct.pack("<4Q", *shape))
        f.write(x.numpy().tobytes())
    rr = subprocess.run([REF, "read", prefix, io_path, out_path], capture_output=True, text=True, timeout=600)
    assert rr.returncode == 0, rr.stdout[-2000:] + rr.stderr[-2000:]
    raw = open(out_path, "rb").read()
    oshape = struct.unpack_from("<4Q", raw, 0)
    y_ref = torch.from_numpy(np.frombuffer(raw, np.float32, int(np.prod(oshape)), 32).copy()).reshape(oshape)
    model = Sequential.from_file(prefix)
    model.set_training(True)
    y = model.forward(x)
    assert tuple(y.shape) == tuple(y_ref.shape)
    assert float((y - y_ref).abs().max()) < 1e-3 * max(1.0, float(y_ref.abs().max()))


def test_device_manager_example_exercises_the_gpu(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "device_manager_example.py")], cwd=str(tmp_path),
                       env=ENV, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Successfully allocated 1073741824 bytes on GPU" in r.stdout and "Round trip over flow 'copies': passed" in r.stdout


COORDINATORS = ["semi_async_pipeline_coordinator.py", "sync_pipeline_coordinator.py", "coordinator_tiny_imagenet.py"]


@pytest.mark.parametrize("script", COORDINATORS)
def test_pipeline_coordinator_on_gpus_over_nccl(tmp_path, script):
    """rank 0 = coordinator, ranks 1..S = stage workers, one GPU per rank where there are enough of them."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    port = 29700 + COORDINATORS.index(script)
    env = dict(ENV, CUDA_MODULE_LOADING="EAGER", BATCH_SIZE="32", SYNTHETIC_TRAIN_SAMPLES="128", SYNTHETIC_TEST_SAMPLES="64")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "examples", script)],
                       cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "Stopped all pipeline stages" in r.stdout and "Validation - Loss" in r.stdout, r.stdout[-2000:]
    assert "nan" not in r.stdout.split("Validation - Loss")[1].splitlines()[0]
