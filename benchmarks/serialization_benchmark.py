#!/usr/bin/env python
"""Wire-codec benchmark (counterpart of benchmarks/serialization_benchmark.cpp: one 16x16x1024x1024-class FP32 tensor;
default here 64 MiB).  Compares the native encoder (libdcnn_host.so) with the pure-Python one."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dcnn_b200.io import binary_serializer as bs  # noqa: E402
from dcnn_b200.parallel.pipeline import CommandType, Job, Message  # noqa: E402

if __name__ == "__main__":
    mib = int(os.environ.get("TENSOR_MIB", 64))
    t = torch.randn(mib * 262144 // (16 * 16), 16, 16, 1)
    msg = Message("stage_1", CommandType.FORWARD_JOB, Job(t, 0), "coordinator")
    for name, fn in (("native", lambda: bs.serialize(msg)), ("python", lambda: bs.serialize_py(msg))):
        fn()
        t0 = time.perf_counter()
        for _ in range(5):
            blob = fn()
        us = (time.perf_counter() - t0) / 5 * 1e6
        print(f"serialize ({name}): {us:.0f} us for {len(blob) / 2**20:.1f} MiB ({len(blob) / us / 1e3:.2f} GB/s)")
    t0 = time.perf_counter()
    for _ in range(5):
        back = bs.deserialize(blob)
    us = (time.perf_counter() - t0) / 5 * 1e6
    print(f"deserialize: {us:.0f} us")
    assert torch.equal(back.payload.data, t)
