"""host:port transport for the message-driven pipeline (ref include/pipeline/tcp_communicator.hpp:113-546,
distributed_coordinator.hpp:26-50, network_stage_worker.hpp:25-115, examples/network_worker.cpp:155-195).

The sockets, io threads, framing and buffer pool are native (csrc/host/tcp_transport.cpp, epoll in place of ASIO); this
module is the message codec on top of it and the three classes a user of the reference expects:

  * ``TcpCommunicator``  -- a ``Communicator`` that listens on a port and dials ``Endpoint.tcp(host, port)`` recipients;
  * ``TcpCoordinator``   -- the reference's DistributedCoordinator: stages are named by host:port, no rendezvous needed;
  * ``TcpStageWorker``   -- ``network_worker <port>``: waits for CONFIG_TRANSFER, serves until SHUTDOWN.

Wire format: the reference's FixedHeader + message header, and its payload types 0-3 (nothing / Job / string / bool)
byte for byte (binary_serializer.hpp:27-176), so control traffic and plain jobs are interoperable.  Two payload types
are added for what the reference leaves unimplemented: 16 = a Job that also carries targets (the last stage owns the
loss), 17 = a JSON document with out-of-line tensors (parameter bundles, status / load reports).  Tensor bytes go from
the tensor's memory to the socket without an intermediate Python copy, and from the receive buffer into the new tensor
with one copy.

On GPUs the fast path between stages stays NCCL / peer memory (``DistCommunicator``, ``PipelineEngine``); this transport
is for CPU stages, for multi-host set-ups without torch.distributed, and as the control plane."""
from __future__ import annotations

import ctypes as C
import dataclasses
import json
import struct
import sys
import threading
import time
import warnings
from typing import List, Optional, Sequence, Tuple

import torch

from ... import _hostlib
from .communicator import Communicator, _restore_tensors, _strip_tensors
from .coordinator import Coordinator, PipelineTimeout
from .protocol import CommandType, Endpoint, Job, LoadTracker, Message
from .stage import PipelineStage

_LE = sys.byteorder == "little"
_E = "<" if _LE else ">"
FIXED_HEADER_SIZE = 10
PT_NONE, PT_JOB, PT_STRING, PT_BOOL, PT_JOB_EXT, PT_DOCUMENT = 0, 1, 2, 3, 16, 17


def _lib():
    l = _hostlib.lib()
    if not getattr(l, "_tcp_ready", False):
        l.dcnn_tcp_create.restype = C.c_void_p
        l.dcnn_tcp_create_error.restype = C.c_char_p
        l._tcp_ready = True
    return l


# ---- encoding: a frame body as a list of byte ranges (bytes objects and tensors) -------------------------------
def _u64(v: int) -> bytes:
    return struct.pack(_E + "Q", int(v))


def _tensor_parts(t: torch.Tensor) -> list:
    """Reference tensor record: u64 ndims, u64 dims[], raw float32."""
    t = t.detach().to("cpu", torch.float32).contiguous()
    if t.dim() == 0:          # ndims == 0 means "empty tensor" in the reference's format: a scalar travels as [1]
        t = t.reshape(1)
    return [_u64(t.dim()) + struct.pack(_E + f"{t.dim()}Q", *t.shape), t]


def _raw_tensor_parts(t: torch.Tensor) -> list:
    """Typed record of payload type 17: u64 ndims, u64 dims[], u64 nbytes, raw bytes in the tensor's own dtype."""
    t = t.detach().to("cpu").contiguous()
    return [_u64(t.dim()) + struct.pack(_E + f"{t.dim()}Q", *t.shape) + _u64(t.numel() * t.element_size()), t]


def _to_document(obj):
    if isinstance(obj, LoadTracker):
        return {"__load_tracker__": dataclasses.asdict(obj)}
    if isinstance(obj, dict):
        return {str(k): _to_document(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [_to_document(v) for v in obj]
    if isinstance(obj, torch.dtype):
        return str(obj)
    return obj


def _from_document(obj):
    if isinstance(obj, dict):
        if set(obj) == {"__load_tracker__"}:
            return LoadTracker(**obj["__load_tracker__"])
        return {k: _from_document(v) for k, v in obj.items()}
    if isinstance(obj, list):
        return [_from_document(v) for v in obj]
    return obj


def encode_body(msg: Message) -> list:
    """Frame body (everything after the FixedHeader) as a list of ``bytes`` / CPU tensors, in wire order."""
    r, s = msg.recipient_id.encode(), msg.sender_id.encode()
    head = _u64(len(r)) + r + _u64(len(s)) + s + struct.pack(_E + "H", int(msg.command_type))
    p = msg.payload
    if p is None:
        return [head + _u64(PT_NONE)]
    if isinstance(p, bool):
        return [head + _u64(PT_BOOL) + (b"\x01" if p else b"\x00")]
    if isinstance(p, str):
        b = p.encode()
        return [head + _u64(PT_STRING) + _u64(len(b)) + b]
    if isinstance(p, Job):
        if p.targets is None and p.data is not None:
            return [head + _u64(PT_JOB) + _u64(p.micro_batch_id), *_tensor_parts(p.data)]
        mask = (1 if p.data is not None else 0) | (2 if p.targets is not None else 0)
        parts = [head + _u64(PT_JOB_EXT) + _u64(p.micro_batch_id) + bytes([mask])]
        for t in (p.data, p.targets):
            if t is not None:
                parts += _tensor_parts(t)
        return parts
    found: List[torch.Tensor] = []
    skeleton = _to_document(_strip_tensors(p, found))
    doc = json.dumps({"doc": skeleton, "dtypes": [str(t.dtype).replace("torch.", "") for t in found]}).encode()
    parts = [head + _u64(PT_DOCUMENT) + _u64(len(doc)) + doc + _u64(len(found))]
    for t in found:
        parts += _raw_tensor_parts(t)
    return parts


def encode_frame(msg: Message) -> bytes:
    """The complete frame as one bytes object (tests, and peers that are not this library)."""
    chunks = []
    for part in encode_body(msg):
        chunks.append(part if isinstance(part, bytes) else part.numpy().tobytes())
    body = b"".join(chunks)
    return struct.pack(_E + "BBQ", 1, 0 if _LE else 1, len(body)) + body


# ---- decoding -------------------------------------------------------------------------------------------------
class _Reader:
    def __init__(self, mv: memoryview):
        version, endian = mv[0], mv[1]
        if version != 1:
            raise ValueError(f"unsupported protocol version {version}")
        self.e = "<" if endian == 0 else ">"
        self.native = (endian == 0) == _LE
        self.mv, self.off = mv, 2
        length = self.u64()
        if len(mv) < FIXED_HEADER_SIZE + length:
            raise ValueError("truncated message")

    def unpack(self, fmt):
        v = struct.unpack_from(self.e + fmt, self.mv, self.off)
        self.off += struct.calcsize(self.e + fmt)
        return v[0]

    def u64(self) -> int:
        return self.unpack("Q")

    def raw(self, n: int) -> memoryview:
        out = self.mv[self.off:self.off + n]
        if len(out) != n:
            raise ValueError("truncated message")
        self.off += n
        return out

    def string(self) -> str:
        return bytes(self.raw(self.u64())).decode()

    def shape(self) -> Tuple[int, ...]:
        return tuple(self.u64() for _ in range(self.u64()))

    def tensor(self, dtype=torch.float32, typed: bool = False) -> torch.Tensor:
        """Reference record (shape, float32 data) or, with ``typed``, the record of payload type 17 (shape, byte
        count, data in ``dtype``)."""
        shape = self.shape()
        n = 1
        for d in shape:
            n *= d
        if not typed and len(shape) == 0:       # reference record with ndims == 0: its empty tensor
            return torch.empty(0, dtype=dtype)
        size = torch.empty(0, dtype=dtype).element_size()
        if typed and self.u64() != n * size:
            raise ValueError("tensor record size does not match its shape")
        data = self.raw(n * size)
        if n == 0:
            return torch.empty(shape, dtype=dtype)
        with warnings.catch_warnings():     # frames decoded from ``bytes`` are read-only; the clone is what is kept
            warnings.simplefilter("ignore", UserWarning)
            t = torch.frombuffer(data, dtype=dtype, count=n).clone()   # the one copy out of the receive buffer
        if not self.native and size > 1:
            t = t.view(torch.uint8).reshape(n, size).flip(1).reshape(-1).view(dtype)
        return t.reshape(shape)


def decode_frame(frame) -> Message:
    rd = _Reader(memoryview(frame).cast("B"))
    recipient, sender = rd.string(), rd.string()
    cmd = CommandType(rd.unpack("H"))
    pt = rd.u64()
    if pt == PT_NONE:
        payload = None
    elif pt == PT_JOB:
        mb = rd.u64()
        payload = Job(rd.tensor(), mb)
    elif pt == PT_STRING:
        payload = rd.string()
    elif pt == PT_BOOL:
        payload = bool(rd.unpack("B"))
    elif pt == PT_JOB_EXT:
        mb = rd.u64()
        mask = rd.unpack("B")
        data = rd.tensor() if mask & 1 else None
        targets = rd.tensor() if mask & 2 else None
        payload = Job(data, mb, targets)
    elif pt == PT_DOCUMENT:
        doc = json.loads(rd.string())
        n = rd.u64()
        tensors = [rd.tensor(getattr(torch, doc["dtypes"][i]), typed=True) for i in range(n)]
        payload = _restore_tensors(_from_document(doc["doc"]), tensors)
    else:
        raise ValueError("Unsupported payload type in MessageData deserialization")
    return Message(recipient, cmd, payload, sender)


# ---- communicator ---------------------------------------------------------------------------------------------
class TcpCommunicator(Communicator):
    """Listens on ``host:port`` (port 0 = pick a free one, see ``.port``) and sends to recipients registered with
    ``connect(id, Endpoint.tcp(host, port))``.  ``io_threads`` sizes the native epoll pool (the reference's
    ``--io-threads``).  A receiver thread decodes arriving frames into the priority queues of ``Communicator``."""

    def __init__(self, comm_id: str = "", host: str = "", port: int = 0, io_threads: int = 2,
                 connect_timeout: float = 30.0):
        super().__init__(comm_id)
        self._l = _lib()
        self._h = self._l.dcnn_tcp_create(host.encode(), C.c_int(int(port)), C.c_int(int(io_threads)))
        if not self._h:
            raise OSError(f"TcpCommunicator: {self._l.dcnn_tcp_create_error().decode()}")
        self._h = C.c_void_p(self._h)
        self.port = int(self._l.dcnn_tcp_port(self._h))
        self.connect_timeout = connect_timeout
        self.decode_errors = 0
        self._closed = False
        self._send_lock = threading.Lock()
        self._rx = threading.Thread(target=self._receive_loop, name=f"tcp-rx-{comm_id}", daemon=True)
        self._rx.start()

    # -- registry ---------------------------------------------------------------------------------------------
    def connect(self, recipient_id: str, endpoint: Endpoint) -> bool:
        if endpoint.kind != "tcp":
            raise ValueError(f"{self.id}: TcpCommunicator needs tcp endpoints, got '{endpoint.kind}'")
        rc = self._l.dcnn_tcp_connect(self._h, recipient_id.encode(), endpoint.host.encode(), C.c_int(endpoint.port),
                                      C.c_int(int(self.connect_timeout * 1000)))
        if rc != 0:
            print(f"{self.id}: {self.last_error()}")
            return False
        return super().connect(recipient_id, endpoint)

    def disconnect(self, recipient_id: str):
        self._l.dcnn_tcp_disconnect(self._h, recipient_id.encode())
        super().disconnect(recipient_id)

    def peer_alive(self, recipient_id: str) -> bool:
        return bool(self._l.dcnn_tcp_peer_alive(self._h, recipient_id.encode()))

    def last_error(self) -> str:
        buf = C.create_string_buffer(512)
        self._l.dcnn_tcp_last_error(self._h, buf, C.c_int(512))
        return buf.value.decode(errors="replace")

    def stats(self) -> dict:
        out = (C.c_uint64 * 9)()
        self._l.dcnn_tcp_stats(self._h, out)
        names = ("bytes_sent", "bytes_received", "frames_sent", "frames_received", "accepted", "disconnects",
                 "protocol_errors", "pool_hits", "pool_misses")
        d = dict(zip(names, (int(v) for v in out)))
        d["decode_errors"] = self.decode_errors
        return d

    # -- output -----------------------------------------------------------------------------------------------
    def send_message(self, msg: Message):
        if not msg.sender_id:
            msg.sender_id = self.id
        self.get_recipient(msg.recipient_id)            # KeyError for an unknown recipient, like the other backends
        parts = encode_body(msg)
        n = len(parts)
        ptrs, lens = (C.c_void_p * n)(), (C.c_size_t * n)()
        keep = []
        for i, part in enumerate(parts):
            if isinstance(part, bytes):
                b = C.c_char_p(part)
                keep.append(b)
                ptrs[i], lens[i] = C.cast(b, C.c_void_p), len(part)
            else:
                keep.append(part)
                ptrs[i], lens[i] = part.data_ptr(), part.numel() * part.element_size()
        rc = self._l.dcnn_tcp_send(self._h, msg.recipient_id.encode(), ptrs, lens, C.c_int(n))
        if rc != 0:
            raise ConnectionError(f"{self.id}: {self.last_error()}")

    def flush_output_messages(self, timeout: float = 30.0) -> bool:
        return bool(self._l.dcnn_tcp_flush(self._h, C.c_int(int(timeout * 1000))))

    # -- input ------------------------------------------------------------------------------------------------
    def _receive_loop(self):
        data, size, token = C.POINTER(C.c_uint8)(), C.c_size_t(), C.c_uint64()
        while not self._closed:
            rc = self._l.dcnn_tcp_recv(self._h, C.c_int(200), C.byref(data), C.byref(size), C.byref(token))
            if rc < 0:
                return
            if rc == 0:
                continue
            try:
                frame = (C.c_uint8 * size.value).from_address(C.addressof(data.contents))
                msg = decode_frame(frame)
            except Exception as e:          # a frame this codec cannot read must not kill the receiver
                self.decode_errors += 1
                print(f"{self.id}: dropped an undecodable frame ({e})")
                msg = None
            finally:
                self._l.dcnn_tcp_release(self._h, token)
            if msg is not None:
                self.enqueue_input_message(msg)

    def close(self):
        if self._closed:
            return
        self.flush_output_messages(5.0)
        self._closed = True
        self._rx.join(timeout=2.0)
        self._l.dcnn_tcp_destroy(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- coordinator and worker -----------------------------------------------------------------------------------
class PipelineStageLost(PipelineTimeout):
    """A stage's connection closed while the coordinator was waiting for it."""


def parse_host_port(s: str, default_host: str = "127.0.0.1") -> Tuple[str, int]:
    """'host:port' or 'port'."""
    if ":" in s:
        h, p = s.rsplit(":", 1)
        return h or default_host, int(p)
    return default_host, int(s)


class TcpCoordinator(Coordinator):
    """Coordinator whose stages are ``network_worker`` processes named by host:port (ref DistributedCoordinator:
    ``Endpoint::network(host, port)`` per stage, distributed_coordinator.hpp:26-50).  ``advertise_host`` is the address
    the workers use to reach this process (it goes into every StageConfig)."""

    def __init__(self, model, optimizer, workers: Sequence, advertise_host: str = "127.0.0.1", listen_host: str = "",
                 listen_port: int = 0, io_threads: int = 2, stage_devices: Optional[Sequence[str]] = None):
        super().__init__(model, optimizer)
        addrs = [parse_host_port(w) if isinstance(w, str) else (w[0], int(w[1])) for w in workers]
        self.num_stages = len(addrs)
        self.comm = TcpCommunicator("coordinator", listen_host, listen_port, io_threads)
        self.coordinator_endpoint = Endpoint.tcp(advertise_host, self.comm.port)
        devs = stage_devices or ["cpu"] * len(addrs)
        self.remote_endpoints = [Endpoint.tcp(h, p, d) for (h, p), d in zip(addrs, devs)]

    def join(self, ctype: CommandType, expected: int, timeout: float = 60.0, raise_on_timeout: bool = True) -> bool:
        """Like Coordinator.join, but a stage whose connection dropped fails the wait at once (``PipelineStageLost``)
        instead of running into the timeout: the transport notices end-of-stream within one io-loop wake-up."""
        deadline = time.monotonic() + timeout
        while True:
            left = deadline - time.monotonic()
            if self.comm.wait_for_count(ctype, expected, max(0.0, min(0.25, left))):
                break
            lost = [n for n in self.stage_names if n in self.comm._recipients and not self.comm.peer_alive(n)]
            if lost and self.is_deployed and ctype != CommandType.CONFIG_RECEIVED:
                self._drain_errors()
                raise PipelineStageLost(f"lost the connection to {', '.join(lost)} while waiting for "
                                        f"{expected} x {ctype.name} (got {self.comm.message_count(ctype)})")
            if left <= 0:
                break
        return super().join(ctype, expected, 0.0, raise_on_timeout)

    def stop(self):
        super().stop()
        self.comm.flush_output_messages()

    def close(self):
        self.comm.close()


class TcpStageWorker(PipelineStage):
    """``network_worker <port> [--gpu]``: one OS process per stage listening on a port.  The coordinator dials it, sends
    CONFIG_TRANSFER (model + optimizer JSON and the three endpoints), the worker dials its neighbours and the
    coordinator back, then serves FORWARD/BACKWARD/UPDATE jobs until SHUTDOWN."""

    def __init__(self, port: int, host: str = "", use_gpu: bool = False, device_index: int = 0, io_threads: int = 2):
        device = torch.device("cuda", device_index) if use_gpu else torch.device("cpu")
        comm = TcpCommunicator(f"worker_{port}", host, port, io_threads)
        super().__init__(comm, f"worker_{comm.port}", device=device)
        self.port = comm.port

    def serve(self):
        """Blocks until SHUTDOWN, then closes the sockets."""
        try:
            self.start(threaded=False)
        finally:
            self.comm.close()
