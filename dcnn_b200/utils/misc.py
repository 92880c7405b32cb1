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
        if self.exporter is not None:
            self.exporter.observe(rec)
        return rec

    exporter = None          # optional PrometheusExporter (METRICS_PORT)


class PrometheusExporter:
    """Scrape endpoint for the numeric fields that pass through a ``MetricsLogger``: every field becomes the gauge
    ``dcnn_<field>{rank=..., kind=...}`` holding its latest value, plus the counter ``dcnn_records_total``.
    ``METRICS_PORT=<port>`` starts one in the stock training loops (rank r listens on port + r); the endpoint is served
    by prometheus_client's background thread, so it stays answerable while a step is running.  The reference has no
    metrics endpoint (stdout only)."""

    def __init__(self, port: int, addr: str = "0.0.0.0", rank: int = 0):
        import prometheus_client as pc
        self._pc = pc
        self.registry = pc.CollectorRegistry()
        self.rank = rank
        self._gauges = {}
        self._records = pc.Counter("dcnn_records", "metric records logged", ["rank", "kind"], registry=self.registry)
        self._server = pc.start_http_server(int(port), addr=addr, registry=self.registry)
        self.port = int(port)

    def observe(self, rec: dict):
        kind = str(rec.get("kind", "step"))
        self._records.labels(rank=str(self.rank), kind=kind).inc()
        for k, v in rec.items():
            if k in ("rank", "kind") or isinstance(v, bool) or not isinstance(v, (int, float)):
                continue
            name = "dcnn_" + "".join(c if c.isalnum() else "_" for c in k)
            g = self._gauges.get(name)
            if g is None:
                g = self._gauges[name] = self._pc.Gauge(name, f"latest value of '{k}'", ["rank", "kind"],
                                                        registry=self.registry)
            g.labels(rank=str(self.rank), kind=kind).set(float(v))

    def close(self):
        srv = self._server
        if isinstance(srv, tuple):          # prometheus_client >= 0.17 returns (server, thread)
            srv[0].shutdown()
            srv[0].server_close()
