"""Failure handling of the message-driven pipeline (the reference prints a warning and carries on, or
hangs): a stage exception reaches the coordinator as an error, lost messages become timeouts, health / status / barrier /
load commands are answered, delayed messages are tolerated, and load reports drive a re-partition."""
import time

import pytest
import torch

from dcnn_b200 import nn as dnn
from dcnn_b200.parallel.pipeline import CommandType, InProcessCommunicator, InProcessCoordinator, PipelineTimeout
from dcnn_b200.train import split_microbatches
from tests.test_pipeline_inprocess import _data, _model


def _coordinator(stages=2, micro=2):
    c = InProcessCoordinator(_model(), dnn.SGD(0.05, 0.9), ["cpu"] * stages)
    c.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
    c.set_num_microbatches(micro)
    c.initialize()
    assert c.deploy_stages()
    c.start()
    return c


def test_stage_exception_surfaces_at_the_coordinator():
    c = _coordinator()
    try:
        x, t = _data(8)
        bad = [torch.randn(4, 7, 8, 8) for _ in range(2)]          # wrong channel count: stage 0's conv rejects it
        with pytest.raises((RuntimeError, PipelineTimeout)) as ei:
            c.async_process_batch(bad, split_microbatches(t, 2), timeout=5.0)
        assert "stage" in str(ei.value).lower() or isinstance(ei.value, PipelineTimeout)
        assert c.errors, "the stage's ERROR_REPORT was not recorded"
        assert "stage_0" in c.errors[-1]
    finally:
        c.stop()


def test_health_status_barrier_and_load_reports():
    c = _coordinator(stages=3, micro=2)
    try:
        assert c.health_check(timeout=5.0)
        c.barrier(timeout=5.0)
        status = c.request_status_from_all_stages()
        assert len(status) == 3
        x, t = _data(8)
        c.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2))
        c.update_parameters()
        loads = c.report_load()
        assert set(loads) == set(c.stage_names)
        assert all(l.avg_forward_time_ms > 0 and l.avg_backward_time_ms > 0 and l.max_memory_usage_mb > 0
                   for l in loads.values())
        parts = c.balance_load()
        assert len(parts) == 3 and parts[0].start_layer == 0 and parts[-1].end_layer == c.model.layer_size()
        assert all(p.end_layer == q.start_layer for p, q in zip(parts, parts[1:]))
    finally:
        c.stop()


def test_delayed_messages_are_tolerated_and_lost_updates_time_out():
    c = _coordinator()
    x, t = _data(8)
    try:
        InProcessCommunicator.fault_hook = lambda m: 0.05 if m.command_type == CommandType.BACKWARD_JOB else None
        t0 = time.perf_counter()
        c.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2), timeout=20.0)
        assert time.perf_counter() - t0 >= 0.05                               # slower ...
        res = c.update_parameters()
        assert res["loss"] == res["loss"] and res["loss"] > 0                 # ... still complete (loss from the last stage)
        # a stage that never acknowledges UPDATE_PARAMETERS must not block the coordinator forever
        InProcessCommunicator.fault_hook = lambda m: "drop" if m.command_type == CommandType.PARAMETERS_UPDATED else None
        c.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2), timeout=20.0)
        with pytest.raises(PipelineTimeout):
            c.update_parameters(timeout=1.0)
    finally:
        InProcessCommunicator.fault_hook = None
        c.stop()


def test_replace_stage_in_process_middle_of_three():
    """Coordinator.replace_stage with in-process endpoints: the middle stage of three is stopped and replaced; both
    neighbours re-dial, every stage reloads the snapshot, and training continues with finite, decreasing loss."""
    from dcnn_b200.parallel.pipeline import Endpoint, PipelineStage
    c = InProcessCoordinator(_model(), dnn.SGD(0.05, 0.0), ["cpu"] * 3)
    c.seed = 3
    c.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
    c.set_num_microbatches(2)
    c.initialize()
    assert c.deploy_stages()
    c.start()
    try:
        x, t = _data(8)
        xs, ts = split_microbatches(x, 2), split_microbatches(t, 2)
        c.async_process_batch(xs, ts)
        first = c.update_parameters()["loss"]
        assert c.collect_current_parameters()
        snapshot = [p.clone() for p in c.model.parameters()]
        c.stages[1].stop()                                    # the stage thread ends: its queue is never served again
        with pytest.raises(PipelineTimeout):
            c.async_process_batch(xs, ts, timeout=1.0)
        comm = InProcessCommunicator("stage_1b")
        fresh = PipelineStage(comm, "stage_1b", device="cpu")
        fresh.start(threaded=True)
        c.stages[1] = fresh
        assert c.replace_stage(1, Endpoint.in_process(comm), timeout=30.0)
        # every stage holds the snapshot again
        assert c.collect_current_parameters()
        for p, q in zip(c.model.parameters(), snapshot):
            assert torch.equal(p, q)
        losses = []
        for _ in range(3):
            c.async_process_batch(xs, ts)
            losses.append(c.update_parameters()["loss"])
        assert all(l == l for l in losses) and losses[-1] < first
    finally:
        c.stop()
