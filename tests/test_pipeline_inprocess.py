"""Pipeline layer tests on the in-process backend (the reference has no tests for its pipeline at all):
pipeline(S stages, M microbatches) == single-device step, sync and semi-async schedules, checkpoint gather,
control commands, fail-fast timeout with fault injection."""
import pytest
import torch

import dcnn_b200 as dc
from dcnn_b200 import nn as dnn
from dcnn_b200.parallel.pipeline import (CommandType, InProcessCommunicator, InProcessCoordinator, PipelineTimeout,
                                         StageConfig)
from dcnn_b200.partitioner import BalancedPartitioner, NaivePartitioner
from dcnn_b200.train import TrainStep, split_microbatches


def _model():
    m = (dnn.SequentialBuilder("pipe_test").input([3, 8, 8])
         .conv2d(8, 3, 3, 1, 1, 1, 1, False, "c1").batchnorm(1e-5, 0.1, True, "bn1").activation("relu", "r1")
         .basic_residual_block(8, 8, 1, "rb1").maxpool2d(2, 2, 2, 2, 0, 0, "mp")
         .basic_residual_block(8, 16, 2, "rb2")
         .avgpool2d(2, 2, 1, 1, 0, 0, "ap").flatten("f").dense(5, True, "fc").build())
    return m


def _data(n=8):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(n, 3, 8, 8, generator=g)
    t = torch.nn.functional.one_hot(torch.randint(0, 5, (n,), generator=g), 5).float().reshape(n, 5, 1, 1)
    return x, t


def _reference_run(stages, micro, steps):
    """Single-device run with the same per-stage initial weights and the same microbatch arithmetic."""
    full = _model()
    parts = NaivePartitioner().get_partitions(full.get_layers(), stages)
    subs = full.split(parts)
    for i, s in enumerate(subs):
        s.set_seed(77 + 1000003 * parts[i].start_layer)
        s.initialize()
    seq = dc.Sequential("ref")
    for s in subs:
        for l in s.layers:
            seq.add(l)
    opt = dnn.SGD(0.05, 0.9)
    opt.attach(seq.parameters(), seq.gradients())
    step = TrainStep(seq, opt, dnn.SoftmaxCrossEntropyLoss(), num_microbatches=micro)
    x, t = _data()
    losses = [float(step(x, t)[0]) for _ in range(steps)]
    return seq, losses


@pytest.mark.parametrize("stages,micro,mode", [(2, 2, "semi_async"), (3, 4, "semi_async"), (2, 2, "sync")])
def test_pipeline_equals_single_device(stages, micro, mode):
    torch.manual_seed(0)
    coord = InProcessCoordinator(_model(), dnn.SGD(0.05, 0.9), ["cpu"] * stages)
    coord.seed = 77
    coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
    coord.loss_on_last_stage = mode == "semi_async"
    coord.set_num_microbatches(micro)
    coord.initialize()
    assert coord.deploy_stages()
    coord.start()
    x, t = _data()
    losses = []
    for _ in range(3):
        xs, ts = split_microbatches(x, micro), split_microbatches(t, micro)
        if mode == "semi_async":
            coord.async_process_batch(xs, ts)
            losses.append(coord.update_parameters()["loss"])
        else:
            losses.append(coord.sync_process_batch(xs, ts))
            coord.update_parameters()
    assert coord.collect_current_parameters()
    ref, ref_losses = _reference_run(stages, micro, 3)
    for a, b in zip(losses, ref_losses):
        assert abs(a - b) < 1e-4, (losses, ref_losses)
    for p, q in zip(coord.model.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=1e-5, rtol=1e-4), (p - q).abs().max()
    for (k, a), (_, b) in zip(coord.model.buffers().items(), ref.buffers().items()):
        assert torch.allclose(a, b, atol=1e-5), k
    assert coord.health_check()
    coord.barrier()
    assert len(coord.request_status_from_all_stages()) == stages
    loads = coord.report_load()
    assert set(loads) == {f"stage_{i}" for i in range(stages)}
    new_parts = coord.balance_load()
    assert new_parts[0].start_layer == 0 and new_parts[-1].end_layer == coord.model.layer_size()
    coord.print_profiling_on_all_stages()
    coord.clear_profiling_data()
    coord.stop()


def test_checkpoint_gather_and_reload(tmp_path):
    coord = InProcessCoordinator(_model(), dnn.Adam(1e-2), ["cpu", "cpu"])
    coord.seed = 5
    coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
    coord.set_num_microbatches(2)
    coord.initialize()
    coord.deploy_stages()
    coord.start()
    x, t = _data()
    coord.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2))
    coord.update_parameters()
    path = str(tmp_path / "snap" / "pipe_model")
    assert coord.save_checkpoint(path)
    loaded = dc.Sequential.from_file(path)
    for p, q in zip(loaded.parameters(), coord.model.parameters()):
        assert torch.equal(p, q)
    # push (modified) parameters back to the stages and read them again
    for p in coord.model.parameters():
        p.mul_(0.5)
    assert coord.send_all_params()
    before = [p.clone() for p in coord.model.parameters()]
    assert coord.collect_current_parameters()
    for p, q in zip(coord.model.parameters(), before):
        assert torch.equal(p, q)
    coord.stop()


def test_stage_config_json_roundtrip():
    coord = InProcessCoordinator(_model(), dnn.Adam(1e-3), ["cpu", "cpu"])
    coord.set_loss_function(dnn.LogSoftmaxCrossEntropyLoss())
    coord.initialize()
    j = coord.stage_configs[1].dumps()
    back = StageConfig.from_json(j)
    assert back.stage_id == "stage_1" and back.is_last and not back.is_first
    assert back.model_config == coord.stage_configs[1].model_config
    assert back.optimizer_config["type"] == "adam" and back.loss_config["type"] == "logsoftmax_crossentropy"


def test_priority_dequeue_and_fault_injection():
    from dcnn_b200.parallel.pipeline import Endpoint, Message
    a, b = InProcessCommunicator("a"), InProcessCommunicator("b")
    a.connect("b", Endpoint.in_process(b))
    a.send_message(Message("b", CommandType.BACKWARD_JOB, None))
    a.send_message(Message("b", CommandType.UPDATE_PARAMETERS, None))
    a.send_message(Message("b", CommandType.FORWARD_JOB, None))
    order = [b.dequeue_input_message().command_type for _ in range(3)]
    assert order == [CommandType.FORWARD_JOB, CommandType.BACKWARD_JOB, CommandType.UPDATE_PARAMETERS]
    # a dropped microbatch must surface as a timeout, not a hang (the reference would wait forever)
    coord = InProcessCoordinator(_model(), dnn.SGD(0.1), ["cpu", "cpu"])
    coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
    coord.set_num_microbatches(2)
    coord.initialize()
    coord.deploy_stages()
    coord.start()
    InProcessCommunicator.fault_hook = lambda m: "drop" if (m.command_type == CommandType.FORWARD_JOB
                                                            and m.recipient_id == "next_stage"
                                                            and m.payload.micro_batch_id == 1) else None
    try:
        x, t = _data()
        with pytest.raises(PipelineTimeout):
            coord.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2), timeout=1.0)
    finally:
        InProcessCommunicator.fault_hook = None
        coord.stop()


def test_partitioners():
    m = _model()
    layers = m.get_layers()
    parts = NaivePartitioner().get_partitions(layers, 4)
    sizes = [p.end_layer - p.start_layer for p in parts]
    n = len(layers)
    assert sizes == [n // 4 + (1 if i < n % 4 else 0) for i in range(4)]
    bal = BalancedPartitioner(input_shape=[4, 3, 8, 8]).get_partitions(layers, 3)
    assert bal[0].start_layer == 0 and bal[-1].end_layer == n
    assert all(a.end_layer == b.start_layer for a, b in zip(bal, bal[1:]))
    costs = BalancedPartitioner(input_shape=[4, 3, 8, 8]).layer_costs(layers)
    worst_bal = max(sum(costs[p.start_layer:p.end_layer]) for p in bal)
    worst_naive = max(sum(costs[p.start_layer:p.end_layer]) for p in NaivePartitioner().get_partitions(layers, 3))
    assert worst_bal <= worst_naive
