"""Checkpoint / config / wire-format compatibility tests (the reference has none of these).
The expected bytes are produced by restating the reference writer's logic inside the test."""
import json
import struct

import numpy as np
import pytest
import torch

import dcnn_b200 as dc
from dcnn_b200 import nn as dnn
from dcnn_b200.io import binary_serializer as bs
from dcnn_b200.io import compression
from dcnn_b200.models import create_mnist_trainer, create_resnet9_cifar10
from dcnn_b200.parallel.pipeline import CommandType, Job, Message


def _reference_bin_bytes(model):
    """Sequential::save_to_file (sequential.hpp:832-861) + Tensor::save (tensor.hpp:625-635): per parameter,
    size_t shape[4] native-endian then raw float32 NCHW."""
    out = b""
    for layer in model.layers:
        if layer.has_parameters():
            for p in layer.parameters():
                shape = list(p.shape) + [1] * (4 - p.dim())
                out += struct.pack("=4Q", *shape) + p.detach().cpu().contiguous().numpy().astype(np.float32).tobytes()
    return out


@pytest.mark.parametrize("factory", [create_mnist_trainer, create_resnet9_cifar10])
def test_snapshot_is_byte_exact_and_round_trips(tmp_path, factory):
    m = factory()
    m.set_seed(3)
    m.initialize()
    path = str(tmp_path / "snapshots" / "model")
    m.save_to_file(path)
    assert open(path + ".bin", "rb").read() == _reference_bin_bytes(m)
    cfg = json.load(open(path + ".json"))
    assert cfg["name"] == m.name() and cfg["is_training"] is True
    assert [l["type"] for l in cfg["layers"]] == [l.type() for l in m.layers]
    m2 = dc.Sequential.from_file(path)
    assert m2.get_config() == m.get_config()
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    x = torch.randn(2, *[cfg["layers"][0]["parameters"]["in_channels"], 28 if "mnist" in m.name() else 32,
                         28 if "mnist" in m.name() else 32])
    m.set_training(False)
    m2.set_training(False)
    assert torch.allclose(m.forward(x), m2.forward(x), atol=1e-6)


def test_state_sidecar_and_weights_only_load(tmp_path):
    m = create_mnist_trainer()
    m.initialize()
    m.forward(torch.randn(4, 1, 28, 28))           # moves the BN running statistics
    path = str(tmp_path / "m")
    m.save_to_file(path, save_state=True)
    m2 = create_mnist_trainer()
    m2.load_weights_file(path + ".bin")
    for (k, a), (_, b) in zip(m.buffers().items(), m2.buffers().items()):
        assert torch.equal(a, b), k


def test_residual_config_quirks_for_json_compat():
    m = create_resnet9_cifar10()
    cfg = m.get_config()
    rb = [l for l in cfg["layers"] if l["type"] == "residual_block"][0]
    assert isinstance(rb["parameters"]["main_path"], str)             # sub-layers are JSON *strings*
    main = json.loads(rb["parameters"]["main_path"])
    assert [s["name"] for s in main] == ["conv2d_0", "bn0", "activation_2", "conv2d_3", "bn0"]
    assert rb["parameters"]["has_projection"] is False and rb["parameters"]["shortcut_path"] == "[]"
    auto = dnn.SequentialBuilder("t").input([3, 8, 8]).conv2d(4, 3, 3).batchnorm().activation("relu").build()
    assert [l.name for l in auto.layers] == ["conv2d_0", "batchnorm_0", "activation_0"]   # layer_size() is 0 until build()


def test_wire_format_python_equals_native_and_reference_layout():
    t = torch.arange(24, dtype=torch.float32).reshape(2, 3, 2, 2)
    msgs = [Message("stage_1", CommandType.FORWARD_JOB, Job(t, 3), "coordinator"),
            Message("coordinator", CommandType.CONFIG_RECEIVED, True, "stage_0"),
            Message("stage_0", CommandType.CONFIG_TRANSFER, '{"stage_id": "stage_0"}', "coordinator"),
            Message("stage_0", CommandType.TRAIN_MODE, None, "coordinator")]
    for m in msgs:
        a, b = bs.serialize_py(m), bs.serialize(m)
        assert a == b
        back = bs.deserialize(a)
        assert back.recipient_id == m.recipient_id and back.sender_id == m.sender_id
        assert back.command_type == m.command_type
        if isinstance(m.payload, Job):
            assert back.payload.micro_batch_id == 3 and torch.equal(back.payload.data, t)
        else:
            assert back.payload == m.payload
    raw = bs.serialize_py(msgs[0])
    assert raw[0] == 1 and raw[1] == 0                                   # version, little-endian tag
    assert struct.unpack_from("<Q", raw, 2)[0] == len(raw) - 10          # FixedHeader.length excludes itself
    off = 10 + 8 + len("stage_1") + 8 + len("coordinator")
    assert struct.unpack_from("<H", raw, off)[0] == int(CommandType.FORWARD_JOB) == 1
    assert struct.unpack_from("<Q", raw, off + 2)[0] == 1                # variant index of Job<float>
    assert struct.unpack_from("<QQ", raw, off + 10) == (3, 4)            # micro_batch_id, ndims


def test_compression_round_trip():
    data = np.random.default_rng(0).integers(0, 4, 100000, dtype=np.uint8).tobytes()
    for codec in compression.available_codecs():
        blob = compression.compress(data, codec)
        assert compression.decompress(blob) == data
    assert len(compression.compress(data, "zstd")) < len(data)
    assert compression.compress(b"", "zstd")[4:] == b"" and compression.decompress(compression.compress(b"", "lz4hc")) == b""


def test_zstd_frames_are_plain_zstd_like_the_reference_and_lz4hc_is_real():
    """ZstdCompressor writes / reads raw zstd frames (magic 0xFD2FB528, content size in the frame header) as the
    reference's ZSTD_compress / ZSTD_getFrameContentSize path does; corrupt input is rejected; LZ4-HC round-trips."""
    import pytest
    if not compression.ZstdCompressor.available():
        pytest.skip("no zstd library")
    data = (np.arange(50000, dtype=np.float32) % 17).tobytes()
    frame = compression.ZstdCompressor.compress(data, 3)
    assert frame[:4] == b"\x28\xb5\x2f\xfd" and len(frame) < len(data) // 4
    assert compression.ZstdCompressor.decompress(frame) == data
    assert len(compression.ZstdCompressor.compress(data, 19)) <= len(frame)
    with pytest.raises(ValueError):
        compression.ZstdCompressor.decompress(b"not a zstd frame at all")
    with pytest.raises(ValueError):
        compression.ZstdCompressor.decompress(frame[:-5])
    if compression.Lz4hcCompressor.available():
        blob = compression.Lz4hcCompressor.compress(data)
        assert len(blob) < len(data) // 4 and compression.Lz4hcCompressor.decompress(blob) == data
        with pytest.raises(ValueError):
            compression.Lz4hcCompressor.decompress(blob[:20])
    assert "zstd" in compression.available_codecs()
