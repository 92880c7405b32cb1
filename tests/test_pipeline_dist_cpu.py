"""Message-driven pipeline over torch.distributed (gloo on CPU): 1 coordinator process + 2 stage workers must take the
same steps as the in-process backend / a single device."""
import os
import socket

import torch
import torch.distributed as dist
from dcnn_b200.utils.spawn import spawn_bounded

from dcnn_b200 import nn as dnn
from dcnn_b200.train import split_microbatches
from tests.test_pipeline_inprocess import _data, _model, _reference_run


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _proc(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    from dcnn_b200.parallel.pipeline import DistributedCoordinator, NetworkStageWorker
    from dcnn_b200.parallel.pipeline.launch import init_distributed, pipeline_groups
    init_distributed("gloo", use_gpu=False)
    stage_ranks = [1, 2]
    store, control, data = pipeline_groups(stage_ranks, use_gpu=False)
    if rank == 0:
        coord = DistributedCoordinator(_model(), dnn.SGD(0.05, 0.9), store, control, 0, stage_ranks)
        coord.seed = 77
        coord.set_loss_function(dnn.SoftmaxCrossEntropyLoss())
        coord.set_num_microbatches(2)
        coord.initialize()
        assert coord.deploy_stages(timeout=60)
        coord.start()
        x, t = _data()
        losses = []
        for _ in range(3):
            coord.async_process_batch(split_microbatches(x, 2), split_microbatches(t, 2))
            losses.append(coord.update_parameters()["loss"])
        assert coord.collect_current_parameters()
        torch.save({"losses": losses, "params": [p.clone() for p in coord.model.parameters()]}, out)
        coord.stop()
    else:
        w = NetworkStageWorker(rank, store, control, data, stage_ranks, use_gpu=False)
        w.start(threaded=False)
    dist.barrier()
    dist.destroy_process_group()


def test_distributed_coordinator_two_workers(tmp_path):
    out = str(tmp_path / "pipe.pt")
    spawn_bounded(_proc, args=(3, _free_port(), out), nprocs=3, timeout_s=300)
    got = torch.load(out)
    ref, ref_losses = _reference_run(2, 2, 3)
    for a, b in zip(got["losses"], ref_losses):
        assert abs(a - b) < 1e-4
    for p, q in zip(got["params"], ref.parameters()):
        assert torch.allclose(p, q, atol=1e-5, rtol=1e-4)
