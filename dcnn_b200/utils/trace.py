"""Chrome-trace export of a training run's per-unit timeline (the reference has host wall-clock tables only,
no trace export).  Inside ``with TraceRecorder("trace.json"):`` every fusion unit's forward / backward of every
``Sequential`` is recorded -- with CUDA events on the launching stream for GPU models (device time, resolved once at the
end, no synchronisation while recording) or the wall clock for CPU models -- and written in the Trace Event format that
chrome://tracing and Perfetto load.  ``DCNN_TRACE_FILE=<path>`` makes the stock training loops record their first
``DCNN_TRACE_STEPS`` (default 3) steps."""
from __future__ import annotations

import json
import os
import time
from typing import List, Optional

from . import runtime as _rt


class TraceRecorder:
    def __init__(self, path: Optional[str] = None, rank: Optional[int] = None):
        self.path = path
        self.rank = int(os.environ.get("RANK", "0")) if rank is None else rank
        self._cpu: List[tuple] = []        # (name, category, t0_s, t1_s)
        self._gpu: List[tuple] = []        # (name, category, e0, e1, device index)
        self._t_origin = time.perf_counter()
        self._gpu_origin = {}

    # -- recording (called by Sequential._timed) ---------------------------------------------------------------
    def timed(self, device, name, category, fn):
        if device.type == "cuda":
            import torch
            idx = device.index if device.index is not None else torch.cuda.current_device()
            if idx not in self._gpu_origin:
                o = torch.cuda.Event(enable_timing=True)
                o.record()
                self._gpu_origin[idx] = (o, time.perf_counter() - self._t_origin)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn()
            e1.record()
            self._gpu.append((name, category, e0, e1, idx))
            return out
        t0 = time.perf_counter()
        out = fn()
        self._cpu.append((name, category, t0 - self._t_origin, time.perf_counter() - self._t_origin))
        return out

    def mark(self, name: str):
        """Instant event (step boundaries)."""
        t = time.perf_counter() - self._t_origin
        self._cpu.append((name, "mark", t, t))

    # -- context manager ---------------------------------------------------------------------------------------
    def __enter__(self):
        _rt.ACTIVE_TRACE = self
        return self

    def __exit__(self, *exc):
        _rt.ACTIVE_TRACE = None
        if self.path:
            self.save(self.path)
        return False

    # -- export ------------------------------------------------------------------------------------------------
    def events(self) -> List[dict]:
        ev = []
        for name, cat, t0, t1 in self._cpu:
            if cat == "mark":
                ev.append({"name": name, "cat": cat, "ph": "i", "s": "p", "ts": t0 * 1e6, "pid": self.rank, "tid": 0})
            else:
                ev.append({"name": name, "cat": cat, "ph": "X", "ts": t0 * 1e6, "dur": max(0.0, (t1 - t0) * 1e6),
                           "pid": self.rank, "tid": 0})
        if self._gpu:
            import torch
            torch.cuda.synchronize()
            for name, cat, e0, e1, idx in self._gpu:
                origin, host_off = self._gpu_origin[idx]
                start_us = host_off * 1e6 + origin.elapsed_time(e0) * 1e3
                ev.append({"name": name, "cat": cat, "ph": "X", "ts": start_us, "dur": max(0.0, e0.elapsed_time(e1) * 1e3),
                           "pid": self.rank, "tid": 1 + idx})
        return sorted(ev, key=lambda e: e["ts"])

    def save(self, path: str):
        os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
        meta = [{"name": "process_name", "ph": "M", "pid": self.rank, "args": {"name": f"rank {self.rank}"}},
                {"name": "thread_name", "ph": "M", "pid": self.rank, "tid": 0, "args": {"name": "host"}}]
        for idx in sorted(self._gpu_origin):
            meta.append({"name": "thread_name", "ph": "M", "pid": self.rank, "tid": 1 + idx,
                         "args": {"name": f"cuda:{idx} (device time)"}})
        with open(path, "w") as f:
            json.dump({"traceEvents": meta + self.events(), "displayTimeUnit": "ms"}, f)
        return path
