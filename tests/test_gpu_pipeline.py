"""SPMD PipelineEngine on real GPUs (>= 2): stages over NCCL per-link communicators == the single-device reference, in
FP32-exact mode (CUDA-core GEMMs) so the comparison is tight; also with per-op CUDA graphs (DCNN_PIPE_GRAPH=1).
Every multi-process launch is bounded (spawn_bounded kills ranks that outlive the deadline)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist

from dcnn_b200 import nn as dnn
from dcnn_b200.partitioner import NaivePartitioner
from dcnn_b200.utils.spawn import spawn_bounded
from tests.test_pipeline_inprocess import _data, _model, _reference_run

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _proc(rank, world, port, out, micro, schedule, graphs):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), DCNN_DISABLE_TC="1",
                      CUDA_MODULE_LOADING="EAGER", DCNN_PIPE_GRAPH="1" if graphs else "0")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    from dcnn_b200.parallel.pipeline import PipelineEngine, make_pipeline_groups
    stage, rep, links, dpg = make_pipeline_groups(world, 1)
    full = _model()
    parts = NaivePartitioner().get_partitions(full.get_layers(), world)
    eng = PipelineEngine(full, parts, stage, dnn.SGD(0.05, 0.9), dnn.SoftmaxCrossEntropyLoss(), micro, links,
                         f"cuda:{rank}", schedule=schedule, seed=77)
    x, t = _data(8)
    losses = []
    for _ in range(5):                       # graphs are captured on the third step: run past it
        loss, _ = eng.train_step(x, t)
        if loss is not None:
            losses.append(float(loss))
    torch.cuda.synchronize()
    gathered = [None] * world
    dist.all_gather_object(gathered, (stage, [p.detach().cpu().contiguous() for p in eng.model.parameters()], losses))
    if rank == 0:
        torch.save(gathered, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("schedule,graphs", [("semi_async", False), ("sync", False), ("semi_async", True)])
def test_engine_on_gpus_equals_single_device(tmp_path, schedule, graphs):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    out = str(tmp_path / "eng.pt")
    spawn_bounded(_proc, args=(2, _free_port(), out, 2, schedule, graphs), nprocs=2, timeout_s=240)
    got = torch.load(out)
    ref, ref_losses = _reference_run(2, 2, 5)
    params = []
    for s in range(2):
        params.extend([g for g in got if g[0] == s][0][1])
    for p, q in zip(params, ref.parameters()):
        qq = q if q.is_contiguous() else q.contiguous()
        assert torch.allclose(p.reshape(qq.shape), qq, atol=2e-4, rtol=2e-3), float((p.reshape(qq.shape) - qq).abs().max())
    losses = [g for g in got if g[0] == 1][0][2]
    for a, b in zip(losses, ref_losses):
        assert abs(a - b) < 1e-3
