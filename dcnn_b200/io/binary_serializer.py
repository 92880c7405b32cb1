"""Wire codec compatible with the reference's BinarySerializer / TBuffer
(/root/reference/include/pipeline/binary_serializer.hpp:27-176, message.hpp:23-37, tbuffer.hpp:184-215):

    FixedHeader : u8 version(=1), u8 endianness (0 little / 1 big), u64 length of the rest
    header      : u64 len + recipient bytes, u64 len + sender bytes, u16 command
    data        : u64 payload_type, then
                  0 -> nothing            1 -> Job  {u64 micro_batch_id, tensor}
                  2 -> string {u64 len, bytes}       3 -> bool {u8}
    tensor      : u64 ndims, u64 dims[ndims], raw float32 (host order)

Only control messages and checkpoints use it -- on the data plane tensors never leave the GPUs.  ``serialize`` uses the
native encoder (csrc/host/codec.cpp) when available; the pure-Python implementation is the specification and the two
are tested to be byte-identical."""
from __future__ import annotations

import ctypes as C
import struct
import sys

import numpy as np
import torch

from .. import _hostlib
from ..parallel.pipeline.protocol import CommandType, Job, Message

_LE = sys.byteorder == "little"
_E = "<" if _LE else ">"
PROTOCOL_VERSION = 1
FIXED_HEADER_SIZE = 10


def _payload_type(payload) -> int:
    if payload is None:
        return 0
    if isinstance(payload, Job):
        return 1
    if isinstance(payload, str):
        return 2
    if isinstance(payload, bool):
        return 3
    raise TypeError(f"Unsupported payload type in MessageData: {type(payload).__name__}")


def _wire_tensor(t: torch.Tensor) -> torch.Tensor:
    """float32 CPU tensor as it goes on the wire.  In the reference's format ``ndims == 0`` is the EMPTY tensor (a
    default-constructed Tensor, binary_serializer.hpp:96-107), so a 0-dim torch scalar travels as shape [1]."""
    t = t.detach().to("cpu", torch.float32).contiguous()
    return t.reshape(1) if t.dim() == 0 else t


def serialize_tensor(t: torch.Tensor) -> bytes:
    a = _wire_tensor(t).numpy()
    return struct.pack(_E + "Q", a.ndim) + struct.pack(_E + f"{a.ndim}Q", *a.shape) + a.tobytes()


def serialize_py(msg: Message) -> bytes:
    r, s = msg.recipient_id.encode(), msg.sender_id.encode()
    body = struct.pack(_E + "Q", len(r)) + r + struct.pack(_E + "Q", len(s)) + s + struct.pack(_E + "H", int(msg.command_type))
    pt = _payload_type(msg.payload)
    body += struct.pack(_E + "Q", pt)
    if pt == 1:
        body += struct.pack(_E + "Q", int(msg.payload.micro_batch_id)) + serialize_tensor(msg.payload.data)
    elif pt == 2:
        b = msg.payload.encode()
        body += struct.pack(_E + "Q", len(b)) + b
    elif pt == 3:
        body += struct.pack("B", 1 if msg.payload else 0)
    return struct.pack(_E + "BBQ", PROTOCOL_VERSION, 0 if _LE else 1, len(body)) + body


def serialize(msg: Message, use_native: bool = True) -> bytes:
    if not (use_native and _hostlib.available()):
        return serialize_py(msg)
    l = _hostlib.lib()
    pt = _payload_type(msg.payload)
    dims = (C.c_uint64 * 8)()
    ndims, tptr, sptr, slen, flag, mb = 0, None, None, 0, 0, 0
    keep = None
    if pt == 1:
        keep = _wire_tensor(msg.payload.data)
        ndims = keep.dim()
        for i, d in enumerate(keep.shape):
            dims[i] = d
        tptr, mb = C.c_void_p(keep.data_ptr()), int(msg.payload.micro_batch_id)
    elif pt == 2:
        keep = msg.payload.encode()
        sptr, slen = keep, len(keep)
    elif pt == 3:
        flag = 1 if msg.payload else 0
    args = (msg.recipient_id.encode(), msg.sender_id.encode(), C.c_uint16(int(msg.command_type)), C.c_uint64(pt),
            C.c_uint64(mb), dims, C.c_uint64(ndims), tptr, sptr, C.c_uint64(slen), C.c_uint8(flag))
    size = l.dcnn_encode_message(None, C.c_size_t(0), *args)
    buf = (C.c_uint8 * size)()
    n = l.dcnn_encode_message(buf, C.c_size_t(size), *args)
    if n != size:
        raise RuntimeError("native message encoder failed")
    return bytes(buf)


def deserialize(buf: bytes) -> Message:
    version, endian, length = struct.unpack_from(_E + "BBQ", buf, 0)
    if version != PROTOCOL_VERSION:
        raise ValueError(f"unsupported protocol version {version}")
    e = "<" if endian == 0 else ">"
    if endian != (0 if _LE else 1):
        (length,) = struct.unpack_from(e + "Q", buf, 2)
    off = FIXED_HEADER_SIZE
    if len(buf) < off + length:
        raise ValueError("truncated message")

    def rd(fmt):
        nonlocal off
        v = struct.unpack_from(e + fmt, buf, off)
        off += struct.calcsize(e + fmt)
        return v[0] if len(v) == 1 else v

    def rstr():
        nonlocal off
        n = rd("Q")
        s = buf[off:off + n].decode()
        off += n
        return s
    recipient, sender = rstr(), rstr()
    cmd = CommandType(rd("H"))
    pt = rd("Q")
    payload = None
    if pt == 1:
        mb = rd("Q")
        nd = rd("Q")
        shape = [rd("Q") for _ in range(nd)]
        n = int(np.prod(shape)) if nd else 0          # ndims == 0: the reference's empty tensor
        arr = np.frombuffer(buf, dtype=np.dtype("float32").newbyteorder(e), count=n, offset=off).astype(np.float32)
        off += 4 * n
        payload = Job(torch.from_numpy(arr.reshape(shape).copy()) if nd else torch.empty(0), mb)
    elif pt == 2:
        payload = rstr()
    elif pt == 3:
        payload = bool(rd("B"))
    elif pt != 0:
        raise ValueError("Unsupported payload type in MessageData deserialization")
    return Message(recipient, cmd, payload, sender)
