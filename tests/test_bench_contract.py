"""bench.py's output contract, exercised through its CPU dry-run mode (BENCH_DEVICE=cpu: same script, same code paths
for model selection, parallelism set-up, the two timed arms, launch accounting and the JSON line; CUDA events, pinned
memory and NCCL are swapped for their CPU stand-ins, and the numbers are marked as meaningless)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
        "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "impl"}


def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _bench(args, nproc=1, port=None, timeout=600):
    port = port or _free_port()      # fixed ports collided with other multi-process tests under pytest-xdist
    cmd = [sys.executable]
    if nproc > 1:
        cmd += ["-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
                "--master-port", str(port)]
    cmd += [os.path.join(ROOT, "bench.py")] + args
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT,
                       env=dict(os.environ, BENCH_DEVICE="cpu", OMP_NUM_THREADS="2"))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, "rank 0 must print exactly ONE JSON line"
    return json.loads(lines[0])


def test_single_process_contract():
    out = _bench(["--steps", "2", "--warmup", "1", "--batch", "4", "--model", "mnist_cnn"])
    assert KEYS <= set(out)
    assert out["n_gpus"] == 1 and out["steps"] == 2 and out["warmup"] >= 3 and out["higher_is_better"] is True
    assert out["scaling"] == "weak" and out["data"] == "synthetic" and out["impl"] == "ours"
    assert out["value"] > 0 and abs(out["value"] - 4 * 1000.0 / out["ms_per_step"]) / out["value"] < 0.02
    cfg = out["config"]
    assert cfg["model"] == "mnist_cnn" and cfg["global_batch"] == 4 and cfg["parallelism"] == "single"
    e2e = out["e2e"]
    assert e2e["value"] > 0 and e2e["h2d_bytes_per_step"] == 4 * (28 * 28 + 10) * 4 and e2e["d2h_bytes_per_step"] == 8
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(out["clocks"])
    assert "DRY RUN" in out["dtype"]                      # a CPU run can never pass for a measurement


def test_headline_defaults_name_the_baseline_config():
    out = _bench(["--steps", "1", "--warmup", "1", "--batch", "2"])
    assert out["metric"] == "train images/sec ResNet-50 64x64 (Tiny-ImageNet shape), max over ranks"
    cfg = out["config"]
    assert cfg["model"] == "ResNet-50-Tiny-ImageNet" and cfg["image"] == "3x64x64" and cfg["classes"] == 200
    assert cfg["optimizer"] == "Adam(lr 1e-4, eps 1e-3, wd 1e-3)" and cfg["loss"] == "LogSoftmaxCrossEntropy"


@pytest.mark.parametrize("par,extra", [("dp2", {"dp_replicas": 2, "pipeline_stages": 1}),
                                       ("pp2", {"dp_replicas": 1, "pipeline_stages": 2})])
def test_two_rank_modes(par, extra):
    out = _bench(["--gpus", "2", "--steps", "2", "--warmup", "1", "--batch", "4", "--model", "mnist_cnn",
                  "--parallelism", par], nproc=2)
    cfg = out["config"]
    assert out["n_gpus"] == 2 and cfg["parallelism"] == par
    assert all(cfg[k] == v for k, v in extra.items())
    assert cfg["global_batch"] == (8 if par == "dp2" else 8)      # dp: 2 x 4; pp: the replica is fed batch x stages
    if par == "pp2":
        assert 0.0 <= out["pipeline_bubble_fraction"] < 1.0


def test_reference_arm_answers_for_unsupported_gpu_counts():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert r.returncode == 0
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert out["impl"] == "reference" and "unavailable" in out


def test_cpu_device_mode_measures_baseline_config_1_on_both_arms():
    """`bench.py --device cpu --model mnist_cnn`: a real CPU measurement (native kernels), labelled as such, and the
    reference arm for the same configuration when the reference subset build is present."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--device", "cpu", "--model", "mnist_cnn",
                        "--batch", "16", "--steps", "3", "--warmup", "3"], capture_output=True, text=True, timeout=600,
                       cwd=ROOT, env={k: v for k, v in os.environ.items() if k not in ("BENCH_DEVICE", "DCNN_CPU_BACKEND")})
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert KEYS <= set(out) and out["value"] > 0
    assert "native kernels" in out["dtype"] and "DRY RUN" not in out["dtype"] and out["config"]["model"] == "mnist_cnn"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--device", "cpu",
                        "--model", "mnist_cnn", "--batch", "16", "--steps", "3", "--warmup", "1"], capture_output=True,
                       text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0
    ref = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert ref["impl"] == "reference" and ("unavailable" in ref or ref["value"] > 0)
