"""Optional compression hook for checkpoint files and control messages (ref include/pipeline/compression_impl/*:
Zstd implemented, LZ4HC/Blosc pass-through stubs, wired into no transport).  On NVLink the links are faster than any
compressor, so nothing on the data plane is ever compressed; this exists for snapshots on disk.
``zstd`` is used when the `zstandard` module is installed, otherwise zlib; ``none`` is a pass-through."""
from __future__ import annotations

import zlib

try:  # pragma: no cover - not installed in the build image
    import zstandard as _zstd
except ImportError:
    _zstd = None

_MAGIC = {b"DCZ0": "none", b"DCZ1": "zlib", b"DCZ2": "zstd"}


def available_codecs():
    return ["none", "zlib"] + (["zstd"] if _zstd else [])


def compress(data: bytes, codec: str = "zstd", level: int = 3) -> bytes:
    if codec == "zstd" and _zstd is None:
        codec = "zlib"
    if codec == "none":
        return b"DCZ0" + data
    if codec == "zlib":
        return b"DCZ1" + zlib.compress(data, level)
    if codec == "zstd":
        return b"DCZ2" + _zstd.ZstdCompressor(level=level).compress(data)
    raise ValueError(f"unknown codec {codec}")


def decompress(blob: bytes) -> bytes:
    tag, body = blob[:4], blob[4:]
    codec = _MAGIC.get(tag)
    if codec == "none":
        return body
    if codec == "zlib":
        return zlib.decompress(body)
    if codec == "zstd":
        if _zstd is None:
            raise RuntimeError("zstd-compressed blob but the zstandard module is not installed")
        return _zstd.ZstdDecompressor().decompress(body)
    raise ValueError("not a dcnn_b200 compressed blob")
