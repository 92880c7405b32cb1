#!/usr/bin/env python
"""Pipeline stage worker -- counterpart of the reference's ``network_worker <port> [--gpu] ...``
(examples/network_worker.cpp:155-195).  Two ways to run it:

  * host:port, like the reference (native TCP transport, no rendezvous):

        python examples/network_worker.py 8001 [--gpu] [--io-threads 4]

    and name the workers to the coordinator example with ``--workers host:8001,host:8002``;

  * as a rank of a torch.distributed job (activations travel over NCCL between GPU stages):

        RANK=<1..S> WORLD_SIZE=<S+1> MASTER_ADDR=127.0.0.1 MASTER_PORT=29500 python examples/network_worker.py --gpu

Either way the process waits for CONFIG_TRANSFER from the coordinator, builds its stage and serves
FORWARD/BACKWARD/UPDATE jobs until SHUTDOWN."""
import argparse
import os

os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

from _common import sys  # noqa: F401  (path setup)

import torch

from dcnn_b200.parallel.pipeline import NetworkStageWorker
from dcnn_b200.parallel.pipeline.launch import init_distributed, pipeline_groups
from dcnn_b200.utils.hardware import HardwareInfo, ThreadAffinity


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("port", nargs="?", type=int, default=None,
                    help="listen on this TCP port (host:port mode); omit to join a torch.distributed job")
    ap.add_argument("--host", default="", help="listen address in host:port mode (default: all interfaces)")
    ap.add_argument("--gpu", action="store_true", help="run the stage on a GPU (index = rank-1 unless --device)")
    ap.add_argument("--device", type=int, default=None)
    ap.add_argument("--num-threads", type=int, default=0, help="CPU threads for a CPU stage")
    ap.add_argument("--pin-numa", action="store_true", help="pin this process to the NUMA node of its GPU")
    ap.add_argument("--show-cores", action="store_true")
    ap.add_argument("--ecore", action="store_true", help="pin a CPU stage to the efficiency cores (ref --ecore)")
    ap.add_argument("--max-ecores", type=int, default=0, help="use at most N efficiency cores (ref --max-ecores)")
    ap.add_argument("--io-threads", type=int, default=2,
                    help="io threads of the native TCP transport in host:port mode (the reference sizes its ASIO pool "
                         "with it); unused in torch.distributed mode, where transport threads belong to gloo / NCCL")
    args = ap.parse_args()
    hw = HardwareInfo()
    hw.initialize()
    affinity = ThreadAffinity(hw)
    if args.ecore and not args.gpu:
        ok = affinity.set_efficiency_cores(args.max_ecores)
        print(f"Efficiency-core affinity {'set' if ok else 'NOT set'}")
    if args.show_cores:
        hw.print_info()
        print(f"performance cores: {hw.get_performance_cores()}  efficiency cores: {hw.get_efficiency_cores()}")
        affinity.print_affinity_info()
    if args.port is not None:
        from dcnn_b200.parallel.pipeline import TcpStageWorker
        dev = args.device if args.device is not None else 0
        if args.gpu:
            torch.cuda.set_device(dev)
            if args.pin_numa:
                affinity.pin_to_gpu_node(dev)
        elif args.num_threads:
            torch.set_num_threads(args.num_threads)
        worker = TcpStageWorker(args.port, args.host, args.gpu, dev, max(1, args.io_threads))
        print(f"Network stage worker listening on port {worker.port} ({'GPU ' + str(dev) if args.gpu else 'CPU'})",
              flush=True)
        worker.serve()
        print(f"worker {worker.port}: shutdown")
        return
    rank, world = init_distributed(use_gpu=args.gpu)
    if rank == 0:
        raise SystemExit("rank 0 is the coordinator; start a coordinator example there")
    dev = args.device if args.device is not None else (rank - 1) % max(1, torch.cuda.device_count())
    if args.gpu:
        torch.cuda.set_device(dev)
        if args.pin_numa:
            affinity.pin_to_gpu_node(dev)
    elif args.num_threads:
        torch.set_num_threads(args.num_threads)
    stage_ranks = list(range(1, world))
    store, control, data = pipeline_groups(stage_ranks, args.gpu)
    print(f"Network stage worker rank {rank} listening ({'GPU ' + str(dev) if args.gpu else 'CPU'})")
    NetworkStageWorker(rank, store, control, data, stage_ranks, args.gpu, dev).start(threaded=False)
    print(f"worker {rank}: shutdown")


if __name__ == "__main__":
    main()
