"""Byte-exact codec for the reference's tensor record (``Tensor::save/load``,
/root/reference/include/tensor/tensor.hpp:625-653):

    size_t shape[4]   (4 x 8 bytes, native endianness; no magic / dtype / version)
    T      data[d0*d1*d2*d3]   (float32 for every shipped model), NCHW row-major

The native fast path (csrc/host/codec.cpp via ctypes) is used for large tensors when the
library is present; the pure-Python path is the definition.
"""
from __future__ import annotations

import struct
from typing import BinaryIO, Sequence

import numpy as np
import torch

_FMT = "=4Q"  # native endianness, LP64 size_t
HEADER_BYTES = 32


def _shape4(shape: Sequence[int]):
    s = [int(d) for d in shape]
    if len(s) > 4:
        raise ValueError("tensor record supports at most 4 dims")
    return s + [1] * (4 - len(s))


def write_tensor(f: BinaryIO, t: torch.Tensor):
    s = _shape4(t.shape)
    f.write(struct.pack(_FMT, *s))
    # logical NCHW order regardless of the in-memory (channels-last) layout
    arr = t.detach().to("cpu", torch.float32).contiguous().numpy()
    f.write(arr.tobytes(order="C"))


def read_tensor(f: BinaryIO) -> torch.Tensor:
    hdr = f.read(HEADER_BYTES)
    if len(hdr) != HEADER_BYTES:
        raise EOFError("unexpected end of weights file (model/config mismatch?)")
    s = struct.unpack(_FMT, hdr)
    n = s[0] * s[1] * s[2] * s[3]
    raw = f.read(4 * n)
    if len(raw) != 4 * n:
        raise EOFError("truncated tensor record")
    arr = np.frombuffer(raw, dtype=np.float32).reshape(s).copy()
    return torch.from_numpy(arr)


def tensor_record_bytes(shape: Sequence[int]) -> int:
    s = _shape4(shape)
    return HEADER_BYTES + 4 * s[0] * s[1] * s[2] * s[3]
