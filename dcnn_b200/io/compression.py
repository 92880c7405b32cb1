"""Compression for checkpoint files and control messages (ref include/pipeline/compression_impl/*: Zstd implemented,
LZ4HC and Blosc pass-through stubs, wired into no transport).  On NVLink the links are faster than any compressor, so
nothing on the data plane is ever compressed; this exists for snapshots on disk and slow inter-host control links.

Zstandard and LZ4 come from the system's ``libzstd.so.1`` / ``liblz4.so.1`` through ctypes (the ``zstandard`` module is
used instead when it is installed); ``zlib`` is the fallback when neither library can be loaded.

Two layers:
  * ``ZstdCompressor`` / ``Lz4hcCompressor`` -- the reference's classes.  ``ZstdCompressor`` emits and reads plain zstd
    frames exactly like the reference's (``ZSTD_compress`` / ``ZSTD_getFrameContentSize`` + ``ZSTD_decompress``), so
    blobs interoperate.  ``Lz4hcCompressor`` is a real LZ4-HC codec here (the reference's is a TODO that returns its
    input): an 8-byte little-endian original size, then one LZ4 block.
  * ``compress`` / ``decompress`` -- tagged blobs (4-byte magic naming the codec) for files written by this library."""
from __future__ import annotations

import ctypes as C
import ctypes.util
import struct
import zlib

try:  # pragma: no cover - not installed in the build image
    import zstandard as _zstd_mod
except ImportError:
    _zstd_mod = None


def _load(names):
    for n in names:
        try:
            return C.CDLL(n)
        except OSError:
            continue
    return None


def _libzstd():
    lib = _load([n for n in ("libzstd.so.1", ctypes.util.find_library("zstd")) if n])
    if lib is None:
        return None
    lib.ZSTD_compressBound.restype = C.c_size_t
    lib.ZSTD_compressBound.argtypes = [C.c_size_t]
    lib.ZSTD_compress.restype = C.c_size_t
    lib.ZSTD_compress.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_int]
    lib.ZSTD_decompress.restype = C.c_size_t
    lib.ZSTD_decompress.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_size_t]
    lib.ZSTD_getFrameContentSize.restype = C.c_ulonglong
    lib.ZSTD_getFrameContentSize.argtypes = [C.c_char_p, C.c_size_t]
    lib.ZSTD_isError.restype = C.c_uint
    lib.ZSTD_isError.argtypes = [C.c_size_t]
    lib.ZSTD_getErrorName.restype = C.c_char_p
    lib.ZSTD_getErrorName.argtypes = [C.c_size_t]
    return lib


def _liblz4():
    lib = _load([n for n in ("liblz4.so.1", ctypes.util.find_library("lz4")) if n])
    if lib is None:
        return None
    lib.LZ4_compressBound.restype = C.c_int
    lib.LZ4_compressBound.argtypes = [C.c_int]
    lib.LZ4_compress_HC.restype = C.c_int
    lib.LZ4_compress_HC.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.LZ4_decompress_safe.restype = C.c_int
    lib.LZ4_decompress_safe.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    return lib


_ZSTD = _libzstd()
_LZ4 = _liblz4()
_CONTENTSIZE_UNKNOWN = 2**64 - 1
_CONTENTSIZE_ERROR = 2**64 - 2


class ZstdCompressor:
    """ref ZstdCompressor (internal_compressor.tpp:10-53): raw zstd frames; empty input passes through."""

    @staticmethod
    def available() -> bool:
        return _ZSTD is not None or _zstd_mod is not None

    @staticmethod
    def compress(data: bytes, compression_level: int = 3) -> bytes:
        if len(data) == 0:
            return data
        if _ZSTD is None:
            if _zstd_mod is None:
                raise RuntimeError("no zstd library available")
            return _zstd_mod.ZstdCompressor(level=compression_level).compress(data)
        bound = _ZSTD.ZSTD_compressBound(len(data))
        out = C.create_string_buffer(bound)
        n = _ZSTD.ZSTD_compress(out, bound, data, len(data), int(compression_level))
        if _ZSTD.ZSTD_isError(n):
            raise RuntimeError("Zstd compression failed: " + _ZSTD.ZSTD_getErrorName(n).decode())
        return out.raw[:n]

    @staticmethod
    def decompress(data: bytes) -> bytes:
        if len(data) == 0:
            return data
        if _ZSTD is None:
            if _zstd_mod is None:
                raise RuntimeError("zstd-compressed data but no zstd library available")
            return _zstd_mod.ZstdDecompressor().decompress(data)
        size = _ZSTD.ZSTD_getFrameContentSize(data, len(data))
        if size == _CONTENTSIZE_ERROR:
            raise ValueError("Invalid zstd compressed data")
        if size == _CONTENTSIZE_UNKNOWN:
            raise ValueError("Cannot determine decompressed size")
        out = C.create_string_buffer(max(1, size))
        n = _ZSTD.ZSTD_decompress(out, size, data, len(data))
        if _ZSTD.ZSTD_isError(n):
            raise ValueError("Zstd decompression failed: " + _ZSTD.ZSTD_getErrorName(n).decode())
        return out.raw[:n]


class Lz4hcCompressor:
    """LZ4-HC block with an 8-byte little-endian original-size prefix (inputs up to 2 GiB - 1, the LZ4 block limit)."""

    @staticmethod
    def available() -> bool:
        return _LZ4 is not None

    @staticmethod
    def compress(data: bytes, compression_level: int = 9) -> bytes:
        if len(data) == 0:
            return data
        if _LZ4 is None:
            raise RuntimeError("liblz4 not available")
        if len(data) > 0x7E000000:
            raise ValueError("LZ4 blocks hold at most 2 113 929 216 bytes")
        bound = _LZ4.LZ4_compressBound(len(data))
        out = C.create_string_buffer(bound)
        n = _LZ4.LZ4_compress_HC(data, out, len(data), bound, int(compression_level))
        if n <= 0:
            raise RuntimeError("LZ4 compression failed")
        return struct.pack("<Q", len(data)) + out.raw[:n]

    @staticmethod
    def decompress(data: bytes) -> bytes:
        if len(data) == 0:
            return data
        if _LZ4 is None:
            raise RuntimeError("liblz4 not available")
        if len(data) < 8:
            raise ValueError("truncated LZ4 blob")
        (size,) = struct.unpack_from("<Q", data, 0)
        if size > 0x7E000000:
            raise ValueError("corrupt LZ4 size prefix")
        out = C.create_string_buffer(max(1, size))
        n = _LZ4.LZ4_decompress_safe(data[8:], out, len(data) - 8, size)
        if n != size:
            raise ValueError("LZ4 decompression failed")
        return out.raw[:n]


_MAGIC = {b"DCZ0": "none", b"DCZ1": "zlib", b"DCZ2": "zstd", b"DCZ3": "lz4hc"}
_TAG = {v: k for k, v in _MAGIC.items()}


def available_codecs():
    return ["none", "zlib"] + (["zstd"] if ZstdCompressor.available() else []) + (["lz4hc"] if _LZ4 is not None else [])


def compress(data: bytes, codec: str = "zstd", level: int = 3) -> bytes:
    """Tagged blob.  A codec whose library is missing falls back to zlib (the tag says what was used)."""
    if codec == "zstd" and not ZstdCompressor.available():
        codec = "zlib"
    if codec == "lz4hc" and _LZ4 is None:
        codec = "zlib"
    if codec == "none":
        return _TAG["none"] + data
    if codec == "zlib":
        return _TAG["zlib"] + zlib.compress(data, min(9, max(0, level)))
    if codec == "zstd":
        return _TAG["zstd"] + ZstdCompressor.compress(data, level)
    if codec == "lz4hc":
        return _TAG["lz4hc"] + Lz4hcCompressor.compress(data, level)
    raise ValueError(f"unknown codec {codec}")


def decompress(blob: bytes) -> bytes:
    tag, body = blob[:4], blob[4:]
    codec = _MAGIC.get(tag)
    if codec == "none":
        return body
    if codec == "zlib":
        return zlib.decompress(body)
    if codec == "zstd":
        return ZstdCompressor.decompress(body)
    if codec == "lz4hc":
        return Lz4hcCompressor.decompress(body)
    raise ValueError("not a dcnn_b200 compressed blob")
