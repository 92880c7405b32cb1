#!/usr/bin/env python
"""Compression benchmark (counterpart of benchmarks/compression_benchmark.cpp: zstd level 3 on a 32x3x1028x1028 FP32
tensor; default here 32x3x256x256).  Uses io/compression.py: zstd and LZ4-HC from the system libraries through ctypes, zlib for comparison."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dcnn_b200.io import compression  # noqa: E402

if __name__ == "__main__":
    hw = int(os.environ.get("HW", 256))
    data = torch.rand(32, 3, hw, hw).mul(255).round().div(255).numpy().tobytes()
    for codec in compression.available_codecs():
        t0 = time.perf_counter()
        blob = compression.compress(data, codec, 3)
        t1 = time.perf_counter()
        out = compression.decompress(blob)
        t2 = time.perf_counter()
        assert out == data
        print(f"{codec:>5}: {len(data) / 2**20:.1f} MiB -> {len(blob) / 2**20:.1f} MiB, compress {t1 - t0:.3f} s, "
              f"decompress {t2 - t1:.3f} s")
