"""Small helpers (ref include/utils/misc.hpp:10-38, memory.hpp:10-22) + structured JSONL metrics."""
from __future__ import annotations

import json
import os
import time
from typing import Callable, Optional


def get_memory_usage_kb() -> int:
    try:
        with open("/proc/self/status") as f:
            for line in f:
                if line.startswith("VmRSS:"):
                    return int(line.split()[1])
    except OSError:
        pass
    return 0


def benchmark(name: str, fn: Callable[[], None], runs: int = 10, sync: Optional[Callable[[], None]] = None) -> float:
    """Average wall time of ``fn`` in ms over ``runs`` (one warm-up call first); prints like the reference helper."""
    fn()
    if sync:
        sync()
    t0 = time.perf_counter()
    for _ in range(runs):
        fn()
    if sync:
        sync()
    ms = (time.perf_counter() - t0) * 1000.0 / runs
    print(f"{name}: {ms:.3f} ms (avg over {runs} runs)")
    return ms


class MetricsLogger:
    """One JSON object per line: step metrics per rank (images/s, step ms, stage busy ms, bubble %, bytes moved)."""

    def __init__(self, path: Optional[str] = None, rank: int = 0):
        self.path, self.rank = path, rank
        if path:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)

    def log(self, **fields):
        rec = {"ts": time.time(), "rank": self.rank, **fields}
        if self.path:
            with open(self.path, "a") as f:
                f.write(json.dumps(rec) + "\n")
        return rec
