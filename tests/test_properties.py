"""Property tests (hypothesis).  Byte-level codecs: whatever message is generated, encode -> decode is the
identity; the native and the pure-Python reference encoders agree byte for byte; truncated or bit-flipped frames are
rejected or decoded without crashing; compression round-trips arbitrary bytes."""
import struct

import pytest
import torch

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import HealthCheck, given, settings  # noqa: E402
from hypothesis import strategies as st  # noqa: E402

from dcnn_b200.io import binary_serializer, compression  # noqa: E402
from dcnn_b200.parallel.pipeline import CommandType, Job, Message  # noqa: E402
from dcnn_b200.parallel.pipeline.tcp import decode_frame, encode_frame  # noqa: E402

COMMANDS = [c for c in CommandType if c not in (CommandType._START, CommandType._COUNT)]
names = st.text(alphabet=st.characters(min_codepoint=32, max_codepoint=0x2FFF, blacklist_categories=("Cs",)), max_size=24)
shapes = st.lists(st.integers(0, 5), min_size=0, max_size=4)


@st.composite
def tensors(draw, dtype=torch.float32):
    shape = draw(shapes)
    n = 1
    for d in shape:
        n *= d
    seed = draw(st.integers(0, 2**31 - 1))
    g = torch.Generator().manual_seed(seed)
    if dtype.is_floating_point:
        return torch.randn(n, generator=g).to(dtype).reshape(shape)
    return torch.randint(-1000, 1000, (n,), generator=g).to(dtype).reshape(shape)


leaf = st.one_of(st.none(), st.booleans(), st.integers(-2**40, 2**40), st.floats(allow_nan=False, allow_infinity=False),
                 names, tensors(), tensors(torch.int64), tensors(torch.float64))
documents = st.recursive(leaf, lambda kids: st.one_of(st.lists(kids, max_size=4),
                                                      st.dictionaries(names.filter(lambda s: s != "__load_tracker__"), kids,
                                                                      max_size=4)), max_leaves=12)
payloads = st.one_of(st.none(), st.booleans(), names,
                     st.builds(Job, tensors(), st.integers(0, 2**40)),
                     st.builds(Job, st.one_of(st.none(), tensors()), st.integers(0, 2**40), st.one_of(st.none(), tensors())),
                     st.dictionaries(names.filter(lambda s: s != "__load_tracker__"), documents, max_size=4),
                     st.lists(documents, min_size=1, max_size=4))
messages = st.builds(Message, names, st.sampled_from(COMMANDS), payloads, names)


def _same(a, b):
    if isinstance(a, torch.Tensor):
        return isinstance(b, torch.Tensor) and a.dtype == b.dtype and tuple(a.shape) == tuple(b.shape) and torch.equal(a, b)
    if isinstance(a, Job):
        # job tensors use the reference's record, where ndims == 0 is the empty tensor: a 0-dim scalar travels as [1]
        def wire(t):
            return t.reshape(1) if isinstance(t, torch.Tensor) and t.dim() == 0 else t
        return isinstance(b, Job) and a.micro_batch_id == b.micro_batch_id and _same(wire(a.data), b.data) \
            and _same(wire(a.targets), b.targets)
    if isinstance(a, dict):
        return isinstance(b, dict) and set(a) == set(b) and all(_same(a[k], b[k]) for k in a)
    if isinstance(a, (list, tuple)):
        return isinstance(b, (list, tuple)) and len(a) == len(b) and all(_same(x, y) for x, y in zip(a, b))
    return type(a) is type(b) and a == b


@settings(max_examples=150, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(messages)
def test_tcp_codec_round_trips_any_message(msg):
    frame = encode_frame(msg)
    assert struct.unpack_from("<Q", frame, 2)[0] == len(frame) - 10
    back = decode_frame(frame)
    assert (back.recipient_id, back.sender_id, back.command_type) == (msg.recipient_id, msg.sender_id, msg.command_type)
    assert _same(msg.payload, back.payload)


@settings(max_examples=100, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(st.builds(Message, names.filter(lambda s: "\x00" not in s), st.sampled_from(COMMANDS),
                 st.one_of(st.none(), st.booleans(), names, st.builds(Job, tensors(), st.integers(0, 2**40))),
                 names.filter(lambda s: "\x00" not in s)))
def test_native_and_python_reference_encoders_agree(msg):
    py = binary_serializer.serialize_py(msg)
    assert binary_serializer.serialize(msg, use_native=True) == py == encode_frame(msg)
    back = binary_serializer.deserialize(py)
    assert back.command_type == msg.command_type and _same(msg.payload, back.payload)


@settings(max_examples=100, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(messages, st.data())
def test_damaged_frames_never_crash_the_decoder(msg, data):
    frame = bytearray(encode_frame(msg))
    cut = data.draw(st.integers(0, len(frame) - 1))
    with pytest.raises((ValueError, struct.error, IndexError, KeyError, TypeError, AttributeError, OverflowError,
                        RuntimeError, UnicodeDecodeError)):
        decode_frame(bytes(frame[:cut]))                     # every proper prefix is rejected
    pos = data.draw(st.integers(10, len(frame) - 1)) if len(frame) > 10 else None
    if pos is not None:
        frame[pos] ^= 1 << data.draw(st.integers(0, 7))
        try:
            decode_frame(bytes(frame))                       # may decode to something else, must not blow up otherwise
        except (ValueError, struct.error, IndexError, KeyError, TypeError, AttributeError, OverflowError, RuntimeError,
                UnicodeDecodeError, MemoryError):
            pass


@settings(max_examples=60, deadline=None)
@given(st.binary(max_size=5000), st.sampled_from(compression.available_codecs()), st.integers(1, 9))
def test_compression_round_trips_arbitrary_bytes(data, codec, level):
    assert compression.decompress(compression.compress(data, codec, level)) == data


@settings(max_examples=120, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(st.integers(1, 5), st.integers(1, 4), st.integers(1, 9), st.integers(1, 9), st.integers(1, 5),
       st.integers(1, 5), st.integers(1, 5), st.integers(1, 3), st.integers(1, 3), st.integers(0, 3), st.integers(0, 3),
       st.booleans())
def test_native_blocked_convolution_matches_aten_for_any_geometry(n, c, h, w, oc, kh, kw, sh, sw, ph, pw, bias):
    """csrc/host/cpu_conv.cpp over random (also degenerate) geometries: kernels wider than the image, padding larger
    than the kernel, strides that skip the last columns."""
    import torch.nn.functional as F
    from dcnn_b200.ops import cpu_native
    if h + 2 * ph < kh or w + 2 * pw < kw:
        return
    g = torch.Generator().manual_seed(n * 1000 + h * 31 + w)
    x, wt = torch.randn(n, c, h, w, generator=g), torch.randn(oc, c, kh, kw, generator=g)
    b = torch.randn(oc, generator=g) if bias else None
    y_ref = F.conv2d(x, wt, b, (sh, sw), (ph, pw))
    y = cpu_native.conv2d_forward(x, wt, b, (sh, sw), (ph, pw))
    assert y.shape == y_ref.shape and torch.allclose(y, y_ref, atol=1e-4, rtol=1e-4)
    dy = torch.randn(y_ref.shape, generator=g)
    dx = cpu_native.conv2d_backward_data(dy, wt, x.shape, (sh, sw), (ph, pw))
    assert torch.allclose(dx, torch.nn.grad.conv2d_input(x.shape, wt, dy, (sh, sw), (ph, pw)), atol=1e-4, rtol=1e-4)
    dw = torch.zeros_like(wt)
    cpu_native.conv2d_backward_weight(dy, x, dw, (sh, sw), (ph, pw))
    assert torch.allclose(dw, torch.nn.grad.conv2d_weight(x, wt.shape, dy, (sh, sw), (ph, pw)), atol=1e-3, rtol=1e-3)


@settings(max_examples=120, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(st.integers(1, 3), st.integers(1, 3), st.integers(1, 11), st.integers(1, 11), st.integers(1, 4), st.integers(1, 4),
       st.integers(1, 4), st.integers(1, 4), st.integers(0, 2), st.integers(0, 2), st.booleans())
def test_native_pooling_matches_aten_for_any_geometry(n, c, h, w, kh, kw, sh, sw, ph, pw, with_nan):
    """csrc/host/cpu_layers.cpp max / average pooling (both the bounds-checked and the in-image fast path), indices and
    backward included; a NaN in a window propagates like ATen's."""
    import torch.nn.functional as F
    from dcnn_b200.ops import cpu_native
    ph, pw = min(ph, kh // 2), min(pw, kw // 2)
    if h + 2 * ph < kh or w + 2 * pw < kw:
        return
    g = torch.Generator().manual_seed(h * 100 + w)
    x = torch.randn(n, c, h, w, generator=g)
    if with_nan:
        x[0, 0, h // 2, w // 2] = float("nan")
    y, idx = cpu_native.maxpool_forward(x, (kh, kw), (sh, sw), (ph, pw))
    y_ref, idx_ref = F.max_pool2d(x, (kh, kw), (sh, sw), (ph, pw), return_indices=True)
    assert torch.equal(torch.isnan(y), torch.isnan(y_ref)) and torch.equal(torch.nan_to_num(y), torch.nan_to_num(y_ref))
    assert torch.equal(idx, idx_ref)
    dy = torch.randn(y_ref.shape, generator=g)
    dx = cpu_native.maxpool_backward(dy, idx, x.shape)
    dx_ref = torch.zeros(n, c, h * w).scatter_add_(2, idx_ref.reshape(n, c, -1), dy.reshape(n, c, -1)).reshape(x.shape)
    assert torch.allclose(dx, dx_ref, atol=1e-6)
    if not with_nan:
        ya = cpu_native.avgpool_forward(x, (kh, kw), (sh, sw), (ph, pw))
        assert torch.allclose(ya, F.avg_pool2d(x, (kh, kw), (sh, sw), (ph, pw), count_include_pad=True), atol=1e-5)
