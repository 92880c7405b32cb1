"""Native TCP transport (csrc/host/tcp_transport.cpp) and the host:port pipeline built on it: codec round trips and
byte compatibility with the reference wire format, framing / ordering / large frames / disconnects between two
communicators, and a coordinator + two ``network_worker <port>`` processes taking the same steps as a single device."""
import os
import subprocess
import sys
import time

import pytest
import torch

from dcnn_b200 import nn as dnn
from dcnn_b200.io import binary_serializer
from dcnn_b200.parallel.pipeline import (CommandType, Endpoint, Job, LoadTracker, Message, StageConfig, TcpCommunicator,
                                         TcpCoordinator)
from dcnn_b200.parallel.pipeline.tcp import decode_frame, encode_frame
from dcnn_b200.train import split_microbatches
from tests.test_pipeline_inprocess import _data, _model, _reference_run

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _same(a, b):
    if isinstance(a, torch.Tensor):
        return isinstance(b, torch.Tensor) and a.dtype == b.dtype and a.shape == b.shape and torch.equal(a, b)
    if isinstance(a, dict):
        return isinstance(b, dict) and a.keys() == b.keys() and all(_same(a[k], b[k]) for k in a)
    if isinstance(a, (list, tuple)):
        return len(a) == len(b) and all(_same(x, y) for x, y in zip(a, b))
    return a == b


def test_codec_reference_payloads_are_byte_identical_to_the_reference_serializer():
    msgs = [Message("stage_0", CommandType.TRAIN_MODE, None, "coordinator"),
            Message("coordinator", CommandType.CONFIG_RECEIVED, True, "stage_1"),
            Message("stage_1", CommandType.CONFIG_TRANSFER, '{"stage_id": "stage_1"}', "coordinator"),
            Message("stage_1", CommandType.FORWARD_JOB, Job(torch.randn(2, 3, 4, 5), 7), "stage_0")]
    for m in msgs:
        frame = encode_frame(m)
        assert frame == binary_serializer.serialize_py(m)
        back = decode_frame(frame)
        assert (back.recipient_id, back.sender_id, back.command_type) == (m.recipient_id, m.sender_id, m.command_type)
        if isinstance(m.payload, Job):
            assert back.payload.micro_batch_id == 7 and torch.equal(back.payload.data, m.payload.data)
        else:
            assert back.payload == m.payload and type(back.payload) is type(m.payload)
        # and the reference-format decoder reads our frames
        assert binary_serializer.deserialize(frame).command_type == m.command_type


def test_codec_extended_payloads_round_trip():
    job = Job(torch.randn(4, 3, 2, 2), 3, torch.randn(4, 10, 1, 1))
    back = decode_frame(encode_frame(Message("s", CommandType.FORWARD_JOB, job, "c"))).payload
    assert back.micro_batch_id == 3 and torch.equal(back.data, job.data) and torch.equal(back.targets, job.targets)
    back = decode_frame(encode_frame(Message("s", CommandType.BACKWARD_JOB, Job(None, 5), "c"))).payload
    assert back.data is None and back.targets is None and back.micro_batch_id == 5
    bundle = {"stage": "stage_0", "params": [torch.randn(3, 3), torch.arange(5)],
              "buffers": {"bn.running_mean": torch.zeros(4, dtype=torch.float64), "empty": torch.empty(0, 2)},
              "count": 3, "loss_sum": 0.25, "flag": None}
    back = decode_frame(encode_frame(Message("coordinator", CommandType.PARAMS_TRANSFER, bundle, "stage_0"))).payload
    assert _same(bundle, back)
    lt = LoadTracker(1.5, 2.5, 37.0, 1024.0)
    assert decode_frame(encode_frame(Message("coordinator", CommandType.LOAD_REPORT, lt, "s"))).payload == lt
    assert decode_frame(encode_frame(Message("coordinator", CommandType.STATUS_RESPONSE, 0.5, "s"))).payload == 0.5


def test_codec_reads_the_other_byte_order():
    m = Message("stage_1", CommandType.FORWARD_JOB, Job(torch.arange(6, dtype=torch.float32).reshape(2, 3), 9), "stage_0")
    frame = bytearray(encode_frame(m))
    # rewrite the frame as a big-endian peer would have produced it
    import struct
    r, s = b"stage_1", b"stage_0"
    body = struct.pack(">Q", len(r)) + r + struct.pack(">Q", len(s)) + s + struct.pack(">H", int(m.command_type))
    body += struct.pack(">Q", 1) + struct.pack(">Q", 9) + struct.pack(">Q", 2) + struct.pack(">2Q", 2, 3)
    body += struct.pack(">6f", *range(6))
    swapped = struct.pack(">BBQ", 1, 1, len(body)) + body
    assert len(swapped) == len(frame)
    back = decode_frame(swapped)
    assert back.payload.micro_batch_id == 9 and torch.equal(back.payload.data, m.payload.data)


def test_endpoint_tcp_json_uses_the_reference_schema():
    ep = Endpoint.tcp("10.0.0.7", 8001)
    j = ep.to_json()
    assert j == {"communication_type_": "tcp", "parameters_": {"host": "10.0.0.7", "port": "8001", "device": "cpu"}}
    back = Endpoint.from_json(j)
    assert (back.kind, back.host, back.port) == ("tcp", "10.0.0.7", 8001)
    cfg = StageConfig("stage_0", {"name": "m"}, {"type": "sgd"}, Endpoint.tcp("a", 1), Endpoint.tcp("b", 2),
                      Endpoint.tcp("c", 3), None, True, False)
    again = StageConfig.from_json(cfg.dumps())
    assert again.next_stage_endpoint.host == "a" and again.coordinator_endpoint.port == 3


@pytest.fixture
def pair():
    a, b = TcpCommunicator("a", "127.0.0.1", 0, io_threads=2), TcpCommunicator("b", "127.0.0.1", 0, io_threads=1)
    assert a.connect("b", Endpoint.tcp("127.0.0.1", b.port)) and b.connect("a", Endpoint.tcp("127.0.0.1", a.port))
    yield a, b
    a.close()
    b.close()


def test_two_communicators_exchange_ordered_frames_and_large_tensors(pair):
    a, b = pair
    n = 200
    for i in range(n):
        a.send_message(Message("b", CommandType.FORWARD_JOB, Job(torch.full((3, 5), float(i)), i)))
    big = torch.randn(8, 64, 128, 128)          # 32 MiB: far beyond the socket buffer, exercises the write queue
    a.send_message(Message("b", CommandType.BACKWARD_JOB, Job(big, 1)))
    a.send_message(Message("b", CommandType.UPDATE_PARAMETERS))
    assert b.wait_for_count(CommandType.UPDATE_PARAMETERS, 1, 60.0)
    assert b.message_count(CommandType.FORWARD_JOB) == n
    for i in range(n):                            # same connection => arrival order == send order
        m = b.dequeue_input_message()             # FORWARD_JOB is served before BACKWARD_JOB (enum order)
        assert m.command_type == CommandType.FORWARD_JOB and m.payload.micro_batch_id == i
        assert m.sender_id == "a" and float(m.payload.data[0, 0]) == float(i)
    m = b.dequeue_input_message()
    assert m.command_type == CommandType.BACKWARD_JOB and torch.equal(m.payload.data, big)
    # and the other direction at the same time as a second large frame
    b.send_message(Message("a", CommandType.PARAMS_TRANSFER, {"params": [big[:2].clone()], "stage": "b"}))
    assert a.wait_for_count(CommandType.PARAMS_TRANSFER, 1, 60.0)
    got = a.dequeue_message_by_type(CommandType.PARAMS_TRANSFER).payload
    assert got["stage"] == "b" and torch.equal(got["params"][0], big[:2])
    sa, sb = a.stats(), b.stats()
    assert sa["frames_sent"] == n + 2 and sb["frames_received"] == n + 2 and sb["bytes_received"] == sa["bytes_sent"]
    assert sb["protocol_errors"] == 0 and sb["decode_errors"] == 0 and sb["pool_hits"] > 0


def test_concurrent_senders_keep_frames_whole(pair):
    import threading
    a, b = pair
    per, threads = 50, 4

    def work(k):
        for i in range(per):
            a.send_message(Message("b", CommandType.FORWARD_JOB, Job(torch.full((257, 33), float(k * 1000 + i)), k * 1000 + i)))
    ts = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert b.wait_for_count(CommandType.FORWARD_JOB, per * threads, 60.0)
    seen = {}
    for m in b.dequeue_all_messages_by_type(CommandType.FORWARD_JOB):
        assert bool((m.payload.data == float(m.payload.micro_batch_id)).all())
        seen.setdefault(m.payload.micro_batch_id // 1000, []).append(m.payload.micro_batch_id % 1000)
    assert all(seen[k] == list(range(per)) for k in range(threads))   # per-sender order is kept


def test_disconnect_is_detected_and_garbage_is_rejected(pair):
    import socket
    a, b = pair
    assert a.peer_alive("b")
    with pytest.raises(KeyError):
        a.send_message(Message("nobody", CommandType.HEALTH_CHECK))
    # a client that does not speak the protocol is dropped without disturbing the others
    s = socket.create_connection(("127.0.0.1", b.port))
    s.sendall(b"GET / HTTP/1.1\r\n\r\n")
    deadline = time.time() + 10
    while b.stats()["protocol_errors"] == 0 and time.time() < deadline:
        time.sleep(0.01)
    assert b.stats()["protocol_errors"] == 1
    s.close()
    a.send_message(Message("b", CommandType.HEALTH_CHECK, "still here"))
    assert b.wait_for_count(CommandType.HEALTH_CHECK, 1, 10.0)
    b.close()
    deadline = time.time() + 10
    while a.peer_alive("b") and time.time() < deadline:
        time.sleep(0.01)
    assert not a.peer_alive("b")
    with pytest.raises(ConnectionError):
        a.send_message(Message("b", CommandType.HEALTH_CHECK))
    assert not TcpCommunicator("c", "127.0.0.1", 0, connect_timeout=0.3).connect("x", Endpoint.tcp("127.0.0.1", 1))


def _start_worker(extra=()):
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1", PYTHONUNBUFFERED="1")
    p = subprocess.Popen([sys.executable, os.path.join(ROOT, "examples", "network_worker.py"), "0", "--host",
                          "127.0.0.1", "--num-threads", "1", *extra], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                         text=True, env=env, cwd=os.path.join(ROOT, "examples"))
    deadline = time.time() + 120
    while time.time() < deadline:
        line = p.stdout.readline()
        if "listening on port" in line:
            return p, int(line.split("listening on port")[1].split()[0])
        if line == "" and p.poll() is not None:
            break
    p.kill()
    raise RuntimeError("network_worker did not come up")


def test_tcp_coordinator_with_two_network_worker_processes_matches_single_device(tmp_path):
    workers = [_start_worker(), _start_worker(("--io-threads", "1"))]
    coord = None
    try:
        coord = TcpCoordinator(_model(), dnn.SGD(0.05, 0.9), [("127.0.0.1", port) for _, port in workers])
        coord.seed = 77
        coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
        coord.set_num_microbatches(2)
        coord.initialize()
        assert coord.deploy_stages(timeout=120)
        coord.start()
        x, t = _data()
        losses = []
        for _ in range(3):
            coord.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2))
            losses.append(coord.update_parameters()["loss"])
        assert coord.health_check() and coord.collect_current_parameters()
        loads = coord.report_load()
        assert len(loads) == 2 and all(isinstance(v, LoadTracker) for v in loads.values())
        assert coord.save_checkpoint(str(tmp_path / "tcp_model"))
        coord.stop()
        for p, _ in workers:
            assert p.wait(timeout=60) == 0
    finally:
        for p, _ in workers:
            if p.poll() is None:
                p.kill()
        if coord is not None:
            coord.close()
    ref, ref_losses = _reference_run(2, 2, 3)
    for a, b in zip(losses, ref_losses):
        assert abs(a - b) < 1e-4
    for p, q in zip(coord.model.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=1e-5, rtol=1e-4)


def test_tcp_coordinator_notices_a_killed_worker_long_before_the_timeout():
    from dcnn_b200.parallel.pipeline import PipelineStageLost
    workers = [_start_worker(), _start_worker()]
    coord = None
    try:
        coord = TcpCoordinator(_model(), dnn.SGD(0.05, 0.9), [("127.0.0.1", port) for _, port in workers])
        coord.seed = 77
        coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
        coord.set_num_microbatches(2)
        coord.initialize()
        assert coord.deploy_stages(timeout=120)
        coord.start()
        x, t = _data()
        coord.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2))
        coord.update_parameters()
        workers[1][0].kill()                      # the last stage dies
        workers[1][0].wait(timeout=30)
        t0 = time.time()
        # whichever notices first: the coordinator's own dead connection (PipelineStageLost), a send into it
        # (ConnectionError), or stage 0 reporting that its neighbour is gone (ERROR_REPORT -> RuntimeError)
        with pytest.raises((PipelineStageLost, ConnectionError, RuntimeError)):
            coord.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2), timeout=120.0)
            coord.update_parameters(timeout=120.0)
        assert time.time() - t0 < 30
    finally:
        for p, _ in workers:
            if p.poll() is None:
                p.kill()
        if coord is not None:
            coord.close()


def _plain_sgd_reference(steps):
    """Single device, SGD without momentum (a replaced stage loses its optimizer moments; plain SGD has none)."""
    import dcnn_b200 as dc
    from dcnn_b200.partitioner import NaivePartitioner
    from dcnn_b200.train import TrainStep
    full = _model()
    parts = NaivePartitioner().get_partitions(full.get_layers(), 2)
    subs = full.split(parts)
    for i, s in enumerate(subs):
        s.set_seed(77 + 1000003 * parts[i].start_layer)
        s.initialize()
    seq = dc.Sequential("ref")
    for s in subs:
        for l in s.layers:
            seq.add(l)
    opt = dnn.SGD(0.05, 0.0)
    opt.attach(seq.parameters(), seq.gradients())
    step = TrainStep(seq, opt, dnn.SoftmaxCrossEntropyLoss(), num_microbatches=2)
    x, t = _data()
    return seq, [float(step(x, t)[0]) for _ in range(steps)]


def test_replacing_a_killed_worker_continues_from_the_snapshot_exactly(tmp_path):
    """Elastic recovery: stage 1 is killed after two steps, a new worker process takes its place
    (Coordinator.replace_stage), every stage is reloaded from the coordinator's snapshot, and steps 3-4 come out exactly
    as in an uninterrupted single-device run."""
    workers = [_start_worker(), _start_worker()]
    coord = None
    try:
        coord = TcpCoordinator(_model(), dnn.SGD(0.05, 0.0), [("127.0.0.1", port) for _, port in workers])
        coord.seed = 77
        coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
        coord.set_num_microbatches(2)
        coord.initialize()
        assert coord.deploy_stages(timeout=120)
        coord.start()
        x, t = _data()
        xs, ts = split_microbatches(x, 2), split_microbatches(t, 2)
        losses = []
        for _ in range(2):
            coord.async_process_batch(xs, ts)
            losses.append(coord.update_parameters()["loss"])
        assert coord.collect_current_parameters()            # the snapshot recovery starts from
        workers[1][0].kill()
        workers[1][0].wait(timeout=30)
        with pytest.raises(Exception):
            coord.async_process_batch(xs, ts, timeout=60.0)
            coord.update_parameters(timeout=60.0)
        workers.append(_start_worker())
        assert coord.replace_stage(1, Endpoint.tcp("127.0.0.1", workers[2][1]), timeout=120)
        for _ in range(2):
            coord.async_process_batch(xs, ts)
            losses.append(coord.update_parameters()["loss"])
        assert coord.collect_current_parameters()
        coord.stop()
        assert workers[0][0].wait(timeout=60) == 0 and workers[2][0].wait(timeout=60) == 0
    finally:
        for p, _ in workers:
            if p.poll() is None:
                p.kill()
        if coord is not None:
            coord.close()
    ref, ref_losses = _plain_sgd_reference(4)
    for a, b in zip(losses, ref_losses):
        assert abs(a - b) < 1e-4
    for p, q in zip(coord.model.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=1e-5, rtol=1e-4)


def test_resilient_runner_replaces_a_killed_worker_and_replays(tmp_path):
    """ResilientBatchRunner: snapshots every 2 batches, a worker is killed after batch 3, the runner obtains a new
    worker from its callback, reloads the snapshot (state after batch 2), replays batch 3 and continues -- the six losses
    and the final weights equal an undisturbed single-device run (plain SGD)."""
    from dcnn_b200.parallel.pipeline.train import ResilientBatchRunner
    workers = [_start_worker(), _start_worker()]
    coord = None
    try:
        coord = TcpCoordinator(_model(), dnn.SGD(0.05, 0.0), [("127.0.0.1", port) for _, port in workers])
        coord.seed = 77
        coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
        coord.set_num_microbatches(2)
        coord.initialize()
        assert coord.deploy_stages(timeout=120)
        coord.start()
        spawned = []

        def replacement(stage_index, error):
            proc, port = _start_worker()
            workers.append((proc, port))
            spawned.append(stage_index)
            return Endpoint.tcp("127.0.0.1", port)
        runner = ResilientBatchRunner(coord, replacement, snapshot_every=2, batch_timeout=60.0)
        x, t = _data()
        xs, ts = split_microbatches(x, 2), split_microbatches(t, 2)
        losses = []
        for step in range(6):
            if step == 3:
                workers[1][0].kill()                     # stage 1 dies between batches 3 and 4
                workers[1][0].wait(timeout=30)
            losses.append(runner.run_batch(xs, ts)["loss"])
        assert spawned == [1] and runner.recoveries == 1
        assert coord.collect_current_parameters()
        coord.stop()
    finally:
        for p, _ in workers:
            if p.poll() is None:
                p.kill()
        if coord is not None:
            coord.close()
    ref, ref_losses = _plain_sgd_reference(6)
    for a, b in zip(losses, ref_losses):
        assert abs(a - b) < 1e-4
    for p, q in zip(coord.model.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=1e-5, rtol=1e-4)
