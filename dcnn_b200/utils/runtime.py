"""Run-time guards around a training loop (SURVEY §5.1, §5.3, §5.5 -- none of which the reference has):

* :class:`StepWatchdog` -- fail fast instead of hanging: if no step completes within ``timeout_s`` the watchdog dumps
  every Python thread's stack and ends the process (exit code 124), which is what lets a launcher / torchrun notice a
  rank that is stuck in a collective whose peer died;
* :func:`checkpoint_on_signal` -- SIGTERM / SIGINT save a resumable checkpoint (:func:`dcnn_b200.io.save_training_state`)
  before the process ends;
* :func:`nvtx_range` -- NVTX ranges around layers / phases when ``DCNN_NVTX=1`` (visible in nsys / ncu timelines);
* :func:`metrics_from_env` -- a :class:`MetricsLogger` writing JSONL to ``$METRICS_FILE`` (one object per step/epoch).
"""
from __future__ import annotations

import contextlib
import faulthandler
import os
import signal
import sys
import threading
import time
from typing import Callable, Optional

from .misc import MetricsLogger

NVTX = os.environ.get("DCNN_NVTX", "0") == "1"
ACTIVE_TRACE = None   # a utils.trace.TraceRecorder while one is recording (Sequential._timed feeds it)
_ACTIVE = []      # running watchdogs; kick() from anywhere (TrainStep, the pipeline engine) feeds all of them


def kick():
    for w in _ACTIVE:
        w.kick()


class StepWatchdog:
    """``wd = StepWatchdog(60).start(); ...; wd.kick()`` after every completed step; ``wd.stop()`` at the end.
    ``timeout_s <= 0`` disables it.  ``on_timeout`` (tests) replaces the default dump-and-exit."""

    def __init__(self, timeout_s: float, name: str = "train", on_timeout: Optional[Callable[[], None]] = None):
        self.timeout_s, self.name, self.on_timeout = float(timeout_s), name, on_timeout
        self._last = time.monotonic()
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.fired = False

    def start(self):
        if self.timeout_s > 0 and self._thread is None:
            self._last = time.monotonic()
            self._thread = threading.Thread(target=self._run, name=f"dcnn-watchdog-{self.name}", daemon=True)
            self._thread.start()
            _ACTIVE.append(self)
        return self

    def kick(self):
        self._last = time.monotonic()

    def stop(self):
        self._stop.set()
        if self in _ACTIVE:
            _ACTIVE.remove(self)
        if self._thread is not None:
            self._thread.join(timeout=2)
            self._thread = None

    def _run(self):
        while not self._stop.wait(min(1.0, self.timeout_s / 4)):
            if time.monotonic() - self._last > self.timeout_s:
                self.fired = True
                if self.on_timeout is not None:
                    self.on_timeout()
                    return
                rank = os.environ.get("RANK", "0")
                sys.stderr.write(f"[dcnn watchdog] rank {rank}: no '{self.name}' step finished in {self.timeout_s:.0f} s; "
                                 "thread stacks follow, then the process exits with code 124\n")
                faulthandler.dump_traceback(file=sys.stderr, all_threads=True)
                sys.stderr.flush()
                os._exit(124)

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()
        return False


def watchdog_from_env(name: str = "train") -> StepWatchdog:
    """``STEP_TIMEOUT_S`` (default 0 = off)."""
    return StepWatchdog(float(os.environ.get("STEP_TIMEOUT_S", "0") or 0), name)


@contextlib.contextmanager
def checkpoint_on_signal(path: str, model, optimizer, get_epoch: Callable[[], int] = lambda: 0,
                         signals=(signal.SIGTERM, signal.SIGINT)):
    """Inside the block, the listed signals write ``<path>`` (model + BN statistics + optimizer state + RNG, see
    io/training_state.py) and then exit 128+signum.  Handlers are restored on exit; main thread only."""
    from ..io import save_training_state
    if threading.current_thread() is not threading.main_thread():
        yield
        return

    def handler(signum, _frame):
        try:
            save_training_state(path, model, optimizer, epoch=get_epoch(), extra={"signal": int(signum)})
            sys.stderr.write(f"[dcnn] signal {signum}: checkpoint written to {path}\n")
        finally:
            sys.exit(128 + int(signum))
    old = {s: signal.signal(s, handler) for s in signals}
    try:
        yield
    finally:
        for s, h in old.items():
            signal.signal(s, h)


@contextlib.contextmanager
def nvtx_range(name: str):
    if NVTX:
        import torch
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(name)
            try:
                yield
            finally:
                torch.cuda.nvtx.range_pop()
            return
    yield


def metrics_from_env(rank: Optional[int] = None) -> MetricsLogger:
    r = int(os.environ.get("RANK", "0")) if rank is None else rank
    path = os.environ.get("METRICS_FILE")
    if path and r != 0:
        root, ext = os.path.splitext(path)
        path = f"{root}.rank{r}{ext}"
    m = MetricsLogger(path, r)
    port = os.environ.get("METRICS_PORT")
    if port:                      # Prometheus scrape endpoint next to (or instead of) the JSONL file; rank r -> port + r
        try:
            from .misc import PrometheusExporter
            m.exporter = PrometheusExporter(int(port) + r, rank=r)
        except Exception as e:    # a busy port must not stop a training run
            print(f"[dcnn] METRICS_PORT={port}: exporter not started ({type(e).__name__}: {e})")
    return m
