"""Bring-up shared by the pipeline coordinator examples.

Launch modes:
  * single process  (``python examples/<coordinator>.py``): every stage is a thread of this process
    (InProcessCoordinator), stage i on GPU i when GPUs are visible and DEVICE_TYPE=GPU, else on the CPU;
  * one process per participant (``torchrun --nproc-per-node S+1 examples/<coordinator>.py``): rank 0 is the
    coordinator (owns the data loader, CPU side), ranks 1..S run a NetworkStageWorker on GPU rank-1.  This is the
    counterpart of starting ``network_worker <port> --gpu`` S times plus a coordinator binary in the reference;
  * host:port, exactly like the reference (``PIPELINE_WORKERS=host:8001,host:8002 python examples/<coordinator>.py`` or
    ``--workers host:8001,host:8002``): the stages are ``python examples/network_worker.py <port>`` processes reached
    over the native TCP transport; ``COORDINATOR_HOST`` is the address they use to call back (default 127.0.0.1).  The reference's ``.env`` keys work too:
    ``WORKER1_HOST`` / ``WORKER1_PORT``, ``WORKER2_...``, ``COORDINATOR_HOST`` / ``COORDINATOR_PORT`` (``env-example``)."""
from __future__ import annotations

import os
import sys

os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")   # stage workers must not lazy-load kernels mid-schedule

import torch


from dcnn_b200.parallel.pipeline import DistributedCoordinator, InProcessCoordinator, NetworkStageWorker
from dcnn_b200.parallel.pipeline.launch import init_distributed, pipeline_groups


def make_coordinator(model, optimizer, loss, num_stages, cfg, partitioner=None, seed=1234):
    """Returns (coordinator | None, is_coordinator).  Worker ranks serve until SHUTDOWN and return (None, False)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    use_gpu = cfg.device_type == "GPU"
    workers = os.environ.get("PIPELINE_WORKERS", "")
    for i, a in enumerate(sys.argv):
        if a == "--workers" and i + 1 < len(sys.argv):
            workers = sys.argv[i + 1]
        elif a.startswith("--workers="):
            workers = a.split("=", 1)[1]
    if not workers and os.environ.get("WORKER1_HOST"):
        # the reference's .env keys (env-example): WORKER<i>_HOST / WORKER<i>_PORT, COORDINATOR_HOST / COORDINATOR_PORT
        found, i = [], 1
        while os.environ.get(f"WORKER{i}_HOST"):
            found.append(f"{os.environ[f'WORKER{i}_HOST']}:{os.environ.get(f'WORKER{i}_PORT', str(8000 + i))}")
            i += 1
        workers = ",".join(found[:num_stages] if num_stages else found)
    if workers:
        from dcnn_b200.parallel.pipeline import TcpCoordinator
        addrs = [w for w in workers.split(",") if w]
        coord = TcpCoordinator(model, optimizer, addrs, os.environ.get("COORDINATOR_HOST", "127.0.0.1"),
                               listen_port=int(os.environ.get("COORDINATOR_PORT", "0")),
                               stage_devices=["cuda:0" if use_gpu else "cpu"] * len(addrs))
    elif world == 1:
        ngpu = torch.cuda.device_count() if use_gpu else 0
        devs = [f"cuda:{i % ngpu}" if ngpu else "cpu" for i in range(num_stages)]
        coord = InProcessCoordinator(model, optimizer, devs)
    else:
        rank, world = init_distributed(use_gpu=use_gpu)
        num_stages = world - 1
        stage_ranks = list(range(1, world))
        if rank != 0 and use_gpu:
            torch.cuda.set_device((rank - 1) % torch.cuda.device_count())
        store, control, data = pipeline_groups(stage_ranks, use_gpu)
        if rank != 0:
            NetworkStageWorker(rank, store, control, data, stage_ranks, use_gpu,
                               (rank - 1) % max(1, torch.cuda.device_count())).start(threaded=False)
            return None, False
        devs = [f"cuda:{(r - 1)}" if use_gpu else "cpu" for r in stage_ranks]
        coord = DistributedCoordinator(model, optimizer, store, control, 0, stage_ranks, devs)
    if os.environ.get("COORDINATOR_NUM_THREADS"):          # the reference sizes the coordinator's thread pool with this
        torch.set_num_threads(max(1, int(os.environ["COORDINATOR_NUM_THREADS"])))
    coord.seed = seed
    coord.set_loss_function(loss)
    if partitioner is not None:
        coord.set_partitioner(partitioner)
    coord.set_num_microbatches(cfg.num_microbatches)
    coord.initialize()
    if not coord.deploy_stages():
        raise SystemExit("Failed to deploy stages")
    coord.start()
    return coord, True
