#!/usr/bin/env python
"""SPMD pipeline trainer: one process per (stage x replica), activations GPU->GPU over per-link NCCL communicators,
1F1B ("semi_async") or GPipe ("sync") schedule, optional data-parallel replicas of the whole pipeline.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29500 \\
        examples/spmd_pipeline_trainer.py --model resnet50_tiny_imagenet --num-stages 4 --device-type GPU \\
        --batch-size 256 --num-microbatches 8

NUM_STAGES x DP_REPLICAS must equal the number of processes.  This is the B200-native counterpart of the reference's
coordinator + network_worker deployment (which the message-driven examples reproduce literally): no coordinator process,
every rank runs the same program.  DEVICE_TYPE=CPU runs the same thing over gloo."""
import os

import torch
import torch.distributed as dist
from _common import get_env, loaders, setup
from data_parallel_trainer import DATASETS

from dcnn_b200 import nn
from dcnn_b200.models import MODEL_ZOO
from dcnn_b200.parallel.pipeline import PipelineEngine, make_pipeline_groups
from dcnn_b200.parallel.pipeline.train import train_engine_model
from dcnn_b200.partitioner import BalancedPartitioner, NaivePartitioner


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
    if rank != 0:
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            cfg = setup()
    else:
        cfg = setup()
    S = get_env("NUM_STAGES", world)
    R = get_env("DP_REPLICAS", max(1, world // S))
    if S * R != world:
        raise SystemExit(f"NUM_STAGES ({S}) x DP_REPLICAS ({R}) must equal the number of processes ({world})")
    name = get_env("MODEL", "resnet9_cifar10")
    key = next((k for k in ("cifar100", "cifar10", "tiny_imagenet", "mnist") if k in name), None)
    if name not in MODEL_ZOO or key is None:
        raise SystemExit(f"unknown MODEL '{name}'")
    factory, data_dir, shape, classes = DATASETS[key]
    gpu = cfg.device_type == "GPU"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if gpu:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", rank=rank, world_size=world)   # lazy communicators: see PipelineEngine
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    stage, replica, links, dp_group = make_pipeline_groups(S, R)
    train_loader, test_loader = loaders(factory, data_dir, shape, classes, cfg,
                                        n_train=get_env("SYNTHETIC_TRAIN_SAMPLES", cfg.batch_size * 8 * R),
                                        n_test=get_env("SYNTHETIC_TEST_SAMPLES", cfg.batch_size * 2 * R))
    model = MODEL_ZOO[name]()
    if get_env("PARTITIONER", "balanced") == "naive":
        parts = NaivePartitioner().get_partitions(model.get_layers(), S)
    else:
        parts = BalancedPartitioner(input_shape=[cfg.batch_size, *shape]).get_partitions(model.get_layers(), S)
    if rank == 0:
        print(f"Pipeline: {S} stage(s) x {R} replica(s), {cfg.num_microbatches} microbatches, partitions "
              f"{[(p.start_layer, p.end_layer) for p in parts]}")
    optimizer = nn.Adam(get_env("LR_INITIAL", 0.001), 0.9, 0.999, 1e-8, get_env("WEIGHT_DECAY", 0.0))
    engine = PipelineEngine(model, parts, stage, optimizer, nn.LossFactory.create_logsoftmax_crossentropy(),
                            cfg.num_microbatches, links, torch.device("cuda", local) if gpu else torch.device("cpu"),
                            schedule=get_env("SCHEDULE", "semi_async"), dp_group=dp_group if R > 1 else None, seed=42)
    history = train_engine_model(engine, train_loader, test_loader, cfg, replica, R, snapshot=get_env("SNAPSHOT", True))
    if rank == 0 and history:
        print(f"Final: train loss {history[-1]['train_loss']:.4f}, validation accuracy {history[-1]['val_acc'] * 100:.2f}%")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
