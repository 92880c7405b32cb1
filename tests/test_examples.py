"""The example entrypoints (counterparts of the reference's examples/*.cpp binaries) run end to end on CPU with
synthetic data -- including the multi-process pipeline launch (torchrun, gloo)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, EPOCHS="1", BATCH_SIZE="8", NUM_MICROBATCHES="2", SYNTHETIC="1", SYNTHETIC_TRAIN_SAMPLES="16",
           SYNTHETIC_TEST_SAMPLES="8", DEVICE_TYPE="CPU", NUM_THREADS="2", PROGRESS_PRINT_INTERVAL="1000")


def _run(args, cwd, timeout=600, **env):
    r = subprocess.run(args, cwd=cwd, env=dict(ENV, **env), capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


SIMPLE_EXAMPLES = [
    ("mnist_cnn_trainer.py", "Validation - Loss"),
    ("cifar10_resnet9.py", "Validation - Loss"),
    ("uji_ips_trainer.py", "Validation - Loss"),
    ("sync_pipeline_coordinator.py", "Stopped all pipeline stages"),
    ("hardware_info_example.py", "Available devices"),
    ("device_manager_example.py", "Memory access test passed"),
    ("tiny_imagenet_visual_check.py", "PNG files"),
    ("device_ptr_example.py", "grow-only"),
    ("cuda_vs_avx2_comparison.py", "native AVX2"),
    ("serve_example.py", "Served 20 requests"),
    ("microbatching_test.py", "Microbatching test PASSED"),
    ("torch_trainer.py", "Validation - Loss"),
]


@pytest.fixture(scope="module")
def simple_example_outputs(tmp_path_factory):
    """The independent single-process examples, started four at a time (each is mostly interpreter start-up and import
    time); the parametrised test below only inspects what they printed."""
    from concurrent.futures import ThreadPoolExecutor

    dirs = {s: tmp_path_factory.mktemp(s.replace(".py", "")) for s, _ in SIMPLE_EXAMPLES}    # not thread-safe: up front

    def one(script):
        cwd = dirs[script]
        r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", script)], cwd=str(cwd), env=dict(ENV),
                           capture_output=True, text=True, timeout=900)
        return script, r
    with ThreadPoolExecutor(max_workers=4) as ex:
        return dict(ex.map(one, [s for s, _ in SIMPLE_EXAMPLES]))


@pytest.mark.parametrize("script,expect", SIMPLE_EXAMPLES)
def test_example_runs(simple_example_outputs, script, expect):
    r = simple_example_outputs[script]
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert expect in r.stdout


def test_mnist_snapshot_then_test_binary(tmp_path):
    # a snapshot is written when validation accuracy improves on 0 (like the reference): use enough test samples
    _run([sys.executable, os.path.join(ROOT, "examples", "mnist_cnn_trainer.py")], cwd=str(tmp_path),
         SYNTHETIC_TEST_SAMPLES="128", EPOCHS="2")
    assert os.path.exists(tmp_path / "model_snapshots" / "mnist_cnn_model.bin")
    out = _run([sys.executable, os.path.join(ROOT, "examples", "mnist_cnn_test.py")], cwd=str(tmp_path))
    assert "Test - Loss" in out


def test_multiprocess_pipeline_launch(tmp_path):
    out = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "3",
                "--master-addr", "127.0.0.1", "--master-port", "29641",
                os.path.join(ROOT, "examples", "semi_async_pipeline_coordinator.py")], cwd=str(tmp_path), timeout=900)
    assert "Deployed 2 pipeline stages" in out and "Stopped all pipeline stages" in out


def test_host_port_pipeline_launch(tmp_path):
    """scripts/local_tcp_pipeline.sh: two `network_worker <port>` processes + a coordinator naming them host:port."""
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = _run(["bash", os.path.join(ROOT, "scripts", "local_tcp_pipeline.sh"),
                os.path.join(ROOT, "examples", "semi_async_pipeline_coordinator.py"), "2", str(port)],
               cwd=str(tmp_path), timeout=900, PYTHONPATH=ROOT)
    assert "Deployed 2 pipeline stages" in out and "Stopped all pipeline stages" in out
    assert out.count(": shutdown") == 2


def test_host_port_pipeline_from_the_reference_env_keys(tmp_path):
    """WORKER<i>_HOST / WORKER<i>_PORT / COORDINATOR_PORT from a .env file, as in the reference's env-example."""
    import socket
    import time
    ports = []
    for _ in range(3):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        ports.append(s.getsockname()[1])
        s.close()
    env = dict(ENV, PYTHONPATH=ROOT, PYTHONUNBUFFERED="1")
    workers = [subprocess.Popen([sys.executable, os.path.join(ROOT, "examples", "network_worker.py"), str(p), "--host",
                                 "127.0.0.1"], cwd=str(tmp_path), env=env, stdout=subprocess.PIPE,
                                stderr=subprocess.STDOUT, text=True) for p in ports[:2]]
    try:
        (tmp_path / ".env").write_text(f"WORKER1_HOST=127.0.0.1\nWORKER1_PORT={ports[0]}\nWORKER2_HOST=127.0.0.1\n"
                                       f"WORKER2_PORT={ports[1]}\nCOORDINATOR_HOST=127.0.0.1\nCOORDINATOR_PORT={ports[2]}\n")
        time.sleep(1.0)
        out = _run([sys.executable, os.path.join(ROOT, "examples", "sync_pipeline_coordinator.py")], cwd=str(tmp_path),
                   timeout=600, PYTHONPATH=ROOT)
        assert "Deployed 2 pipeline stages" in out and "Stopped all pipeline stages" in out
        for w in workers:
            assert w.wait(timeout=60) == 0
    finally:
        for w in workers:
            if w.poll() is None:
                w.kill()


def test_streaming_loader_in_the_stock_training_loop(tmp_path):
    out = _run([sys.executable, os.path.join(ROOT, "examples", "cifar10_resnet9.py")], cwd=str(tmp_path),
               PREFETCH_DEPTH="2", EPOCHS="2")
    assert out.count("Validation - Loss") == 2 and "Training batches: 2" in out


def test_flag_overrides_reach_an_example(tmp_path):
    out = _run([sys.executable, os.path.join(ROOT, "examples", "mnist_cnn_trainer.py"), "--epochs", "2",
                "--batch-size=4"], cwd=str(tmp_path))
    assert "Epoch 2/2" in out


def test_data_parallel_trainer_two_replicas(tmp_path):
    out = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                "--master-addr", "127.0.0.1", "--master-port", "29643",
                os.path.join(ROOT, "examples", "data_parallel_trainer.py"), "--model", "mnist_cnn"],
               cwd=str(tmp_path), timeout=900, EPOCHS="2", SYNTHETIC_TRAIN_SAMPLES="64", SYNTHETIC_TEST_SAMPLES="32")
    assert "Data parallel: 2 replica(s), per-replica batch 8, global batch 16" in out
    assert out.count("Validation - Loss") == 2          # replica 0 only
    assert "Training batches: 4" in out and "Final: train loss" in out


@pytest.mark.parametrize("procs,stages", [(2, 2), (4, 2)])
def test_spmd_pipeline_trainer(tmp_path, procs, stages):
    out = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(procs),
                "--master-addr", "127.0.0.1", "--master-port", str(29650 + procs),
                os.path.join(ROOT, "examples", "spmd_pipeline_trainer.py"), "--model", "mnist_cnn",
                "--num-stages", str(stages)], cwd=str(tmp_path), timeout=900, EPOCHS="2",
               SYNTHETIC_TRAIN_SAMPLES="64", SYNTHETIC_TEST_SAMPLES="32", NUM_THREADS="1")
    assert f"Pipeline: {stages} stage(s) x {procs // stages} replica(s)" in out
    assert out.count("Validation - Loss") == 2 and "Final: train loss" in out
    # training loss decreases from epoch 1 to epoch 2 on the 64 synthetic samples
    losses = [float(l.split("Loss:")[1].split(",")[0]) for l in out.splitlines() if l.startswith("Training   - Loss")]
    assert len(losses) == 2 and losses[1] < losses[0]
    if "model saved to" in out:          # written whenever validation accuracy improved on 0
        assert os.path.exists(tmp_path / "model_snapshots" / "mnist_cnn_model.bin")
        assert os.path.exists(tmp_path / "model_snapshots" / "mnist_cnn_model.state.bin")


def test_package_command_line(tmp_path, capsys, monkeypatch):
    """python -m dcnn_b200 info | models | summary | train --snapshot | predict (one real `-m` launch, the rest through
    the same main() in this process)."""
    from dcnn_b200.__main__ import main
    monkeypatch.chdir(tmp_path)
    for k, v in ENV.items():
        monkeypatch.setenv(k, v)
    out = _run([sys.executable, "-m", "dcnn_b200", "info"], cwd=str(tmp_path), PYTHONPATH=ROOT)
    assert "dcnn_b200" in out and "devices: CPU:0" in out

    def run(*a):
        assert main(list(a)) == 0
        return capsys.readouterr().out
    assert "resnet50_tiny_imagenet" in run("models")
    assert "Model Summary" in run("summary", "--model", "resnet9_cifar10", "--batch", "2")
    snap = str(tmp_path / "snap" / "mnist")
    os.makedirs(os.path.dirname(snap))
    out = run("train", "--model", "mnist_cnn", "--synthetic", "--batch-size", "8", "--snapshot", snap)
    assert "Snapshot written" in out and os.path.exists(snap + ".bin")
    out = run("predict", "--snapshot", snap, "--input", "4,1,28,28", "--repeat", "2")
    assert "batchnorm 0" in out and "images/s" in out
