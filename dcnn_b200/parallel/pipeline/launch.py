"""Process-group plumbing for the distributed (one process per participant) pipeline.

Rank layout of the message-driven mode: rank 0 = coordinator (CPU side: owns the data loader), ranks 1..S = stages,
one GPU each.  Control messages and coordinator<->stage tensors use a gloo group over all ranks; stage<->stage
activations/gradients use an NCCL group over the stage ranks (p2p over NVLink 5)."""
from __future__ import annotations

import os
from typing import Optional, Sequence

import torch
import torch.distributed as dist


def init_distributed(backend: Optional[str] = None, use_gpu: Optional[bool] = None):
    """Join the torch.distributed job described by RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT (torchrun)."""
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if use_gpu is None:
        use_gpu = torch.cuda.is_available()
    if backend is None:
        backend = "cpu:gloo,cuda:nccl" if use_gpu else "gloo"
    dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world


def pipeline_groups(stage_ranks: Sequence[int], use_gpu: bool):
    """-> (store, control_group, data_group).  Must be called by every rank."""
    from torch.distributed import distributed_c10d as c10d
    store = dist.PrefixStore("dcnn_pipeline", c10d._get_default_store())
    control = dist.new_group(backend="gloo")
    data = dist.new_group(ranks=list(stage_ranks), backend="nccl" if use_gpu else "gloo")
    return store, control, data
