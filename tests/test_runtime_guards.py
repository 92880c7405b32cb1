"""Watchdog, checkpoint-on-signal, JSONL metrics and NVTX plumbing (SURVEY §5.1/5.3/5.5 'new build' items)."""
import json
import os
import signal
import subprocess
import sys
import textwrap
import time


from dcnn_b200.utils import runtime

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_watchdog_fires_only_without_kicks():
    fired = []
    wd = runtime.StepWatchdog(0.4, on_timeout=lambda: fired.append(1)).start()
    for _ in range(8):
        time.sleep(0.1)
        runtime.kick()
    assert not fired and not wd.fired
    time.sleep(1.0)
    assert fired and wd.fired
    wd.stop()
    assert wd not in runtime._ACTIVE
    off = runtime.StepWatchdog(0).start()          # disabled
    assert off._thread is None
    off.stop()


def test_watchdog_kills_a_stuck_process():
    code = textwrap.dedent(f"""
        import sys, time
        sys.path.insert(0, {ROOT!r})
        from dcnn_b200.utils.runtime import StepWatchdog
        StepWatchdog(0.5, "unit").start()
        time.sleep(30)
    """)
    t0 = time.time()
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=60)
    assert r.returncode == 124 and time.time() - t0 < 25
    assert "no 'unit' step finished" in r.stderr and "Thread" in r.stderr or "File" in r.stderr


def test_checkpoint_on_sigterm_is_resumable(tmp_path):
    path = str(tmp_path / "ck" / "m")
    code = textwrap.dedent(f"""
        import os, signal, sys, time
        sys.path.insert(0, {ROOT!r})
        import torch
        from dcnn_b200 import nn as dnn
        from dcnn_b200.models import create_mnist_trainer
        from dcnn_b200.train import TrainStep
        from dcnn_b200.utils.runtime import checkpoint_on_signal
        m = create_mnist_trainer(); m.set_seed(1); m.initialize()
        o = dnn.Adam(1e-2); o.attach(m.parameters(), m.gradients())
        step = TrainStep(m, o, dnn.SoftmaxCrossEntropyLoss())
        x = torch.randn(4, 1, 28, 28); t = torch.zeros(4, 10, 1, 1); t[:, 3] = 1
        with checkpoint_on_signal({path!r}, m, o, lambda: 7):
            step(x, t); step(x, t)
            print("READY", flush=True)
            os.kill(os.getpid(), signal.SIGTERM)
            time.sleep(30)
    """)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert r.returncode == 128 + signal.SIGTERM, r.stderr[-2000:]
    assert "READY" in r.stdout and "checkpoint written" in r.stderr
    from dcnn_b200 import nn as dnn
    from dcnn_b200.io import load_training_state
    from dcnn_b200.models import create_mnist_trainer
    m = create_mnist_trainer()
    m.set_seed(2)
    m.initialize()
    o = dnn.Adam(1e-2)
    o.attach(m.parameters(), m.gradients())
    blob = load_training_state(path, m, o)
    assert blob["epoch"] == 7 and blob["extra"]["signal"] == signal.SIGTERM and o.t == 2


def test_metrics_jsonl_from_training_loop(tmp_path, monkeypatch):
    from dcnn_b200 import nn as dnn
    from dcnn_b200.data import SyntheticDataLoader
    from dcnn_b200.models import create_mnist_trainer
    from dcnn_b200.train import TrainingConfig, train_classification_model
    mf = tmp_path / "metrics" / "run.jsonl"
    monkeypatch.setenv("METRICS_FILE", str(mf))
    monkeypatch.setenv("STEP_TIMEOUT_S", "120")
    cfg = TrainingConfig()
    cfg.epochs, cfg.batch_size, cfg.snapshot_dir, cfg.progress_print_interval = 2, 8, str(tmp_path / "snap"), 1000
    m = create_mnist_trainer()
    m.set_seed(0)
    m.initialize()
    hist = train_classification_model(m, SyntheticDataLoader(32, (1, 28, 28), 10, seed=0),
                                      SyntheticDataLoader(16, (1, 28, 28), 10, seed=1), dnn.Adam(1e-3),
                                      dnn.SoftmaxCrossEntropyLoss(), cfg)
    recs = [json.loads(l) for l in open(mf)]
    assert len(recs) == 2 == len(hist)
    assert recs[1]["kind"] == "epoch" and recs[1]["epoch"] == 2 and recs[1]["images_per_s"] > 0 and recs[1]["rank"] == 0
    assert not runtime._ACTIVE            # the loop's watchdog was stopped
    assert runtime.metrics_from_env(rank=3).path.endswith("run.rank3.jsonl")


def test_nvtx_range_is_a_noop_without_cuda():
    with runtime.nvtx_range("x"):
        pass


def test_flag_override_layer(tmp_path, monkeypatch):
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import _common
    from dcnn_b200.utils import env
    monkeypatch.chdir(tmp_path)
    (tmp_path / ".env").write_text("EPOCHS=9\nBATCH_SIZE=16\n")
    monkeypatch.setattr(sys, "argv", ["prog", "--epochs", "3", "--device-type=CPU", "positional", "--synthetic"])
    env.clear_loaded()
    try:
        cfg = _common.setup()
        assert cfg.epochs == 3 and cfg.batch_size == 16 and cfg.device_type == "CPU"
        assert env.get_env("SYNTHETIC", False) is True
        assert _common.apply_flag_overrides(["a", "--x-y=1", "b"]) == ["a", "b"] and os.environ["X_Y"] == "1"
    finally:
        env.clear_loaded()
        for k in ("EPOCHS", "BATCH_SIZE", "DEVICE_TYPE", "SYNTHETIC", "X_Y"):
            os.environ.pop(k, None)


def test_chrome_trace_of_training_steps(tmp_path):
    from dcnn_b200 import nn as dnn
    from dcnn_b200.models import create_mnist_trainer
    from dcnn_b200.train import TrainStep
    from dcnn_b200.utils.trace import TraceRecorder
    m = create_mnist_trainer()
    m.set_seed(0)
    m.initialize()
    o = dnn.SGD(0.01)
    o.attach(m.parameters(), m.gradients())
    step = TrainStep(m, o, dnn.SoftmaxCrossEntropyLoss())
    import torch
    x = torch.randn(4, 1, 28, 28)
    t = torch.zeros(4, 10, 1, 1)
    t[:, 1] = 1
    path = str(tmp_path / "t" / "trace.json")
    with TraceRecorder(path) as rec:
        for i in range(2):
            rec.mark(f"step {i}")
            step(x, t)
    assert runtime.ACTIVE_TRACE is None
    blob = json.load(open(path))
    ev = [e for e in blob["traceEvents"] if e.get("ph") == "X"]
    units = len(m._get_units())
    assert len(ev) == 2 * 2 * units                                # 2 steps x (forward + backward) x units
    assert {e["cat"] for e in ev} == {"forward", "backward"} and all(e["dur"] >= 0 for e in ev)
    assert [e["ts"] for e in ev] == sorted(e["ts"] for e in ev)
    assert sum(1 for e in blob["traceEvents"] if e.get("ph") == "i") == 2
    # forward of a step precedes its backward; the last unit's backward comes first in the backward sweep
    first_step = ev[:2 * units]
    assert [e["cat"] for e in first_step] == ["forward"] * units + ["backward"] * units
    assert first_step[units]["name"] == first_step[units - 1]["name"]


def test_trace_file_env_records_first_steps(tmp_path, monkeypatch):
    import torch
    from dcnn_b200 import nn as dnn
    from dcnn_b200.models import create_mnist_trainer
    from dcnn_b200.train import TrainStep
    path = tmp_path / "run.json"
    monkeypatch.setenv("DCNN_TRACE_FILE", str(path))
    monkeypatch.setenv("DCNN_TRACE_STEPS", "2")
    m = create_mnist_trainer()
    m.initialize()
    o = dnn.SGD(0.01)
    o.attach(m.parameters(), m.gradients())
    step = TrainStep(m, o, dnn.SoftmaxCrossEntropyLoss())
    x, t = torch.randn(2, 1, 28, 28), torch.zeros(2, 10, 1, 1)
    t[:, 0] = 1
    for _ in range(4):
        step(x, t)
    assert runtime.ACTIVE_TRACE is None and path.exists()
    ev = [e for e in json.load(open(path))["traceEvents"] if e.get("ph") == "X"]
    assert len(ev) == 2 * 2 * len(m._get_units())            # only the first two steps


def test_prometheus_exporter_serves_logged_metrics(monkeypatch):
    """METRICS_PORT: numeric fields of every MetricsLogger record are scrapeable as dcnn_<field>{rank, kind}."""
    import socket
    import urllib.request
    pytest = __import__("pytest")
    pytest.importorskip("prometheus_client")
    from dcnn_b200.utils.runtime import metrics_from_env
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    monkeypatch.setenv("METRICS_PORT", str(port))
    monkeypatch.delenv("METRICS_FILE", raising=False)
    m = metrics_from_env(rank=0)
    assert m.exporter is not None
    try:
        m.log(kind="epoch", images_per_s=1234.5, train_loss=0.25, note="text is skipped", ok=True)
        m.log(kind="epoch", images_per_s=2000.0, train_loss=0.125)
        body = urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics", timeout=10).read().decode()
        assert 'dcnn_images_per_s{kind="epoch",rank="0"} 2000.0' in body
        assert 'dcnn_train_loss{kind="epoch",rank="0"} 0.125' in body
        assert 'dcnn_records_total{kind="epoch",rank="0"} 2.0' in body
        assert "dcnn_note" not in body and "dcnn_ok" not in body
    finally:
        m.exporter.close()
