#!/usr/bin/env python
"""The fused reduce-scatter + optimizer + all-gather kernel (csrc/comm.cu) in isolation against (a) NCCL all-reduce +
the multi-tensor update and (b) the NVLink roofline.  Launch with torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 \\
        benchmarks/allreduce_roofline.py

Per step each GPU reads (R-1)/R of the gradients from its peers and writes (R-1)/R of the parameters to them; the
roofline is those bytes over the measured unidirectional NVLink bandwidth (MEASURED_PEAKS.json if present, else the
900 GB/s nominal).  Device-timed with CUDA events, max over ranks.  (>= 70 % of the roofline.)"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200 import nn  # noqa: E402
from dcnn_b200.models import create_resnet50_tiny_imagenet  # noqa: E402
from dcnn_b200.parallel.dp import DataParallel  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = f"cuda:{local}"

    def build(baseline, num_blocks=0):
        m = create_resnet50_tiny_imagenet()
        m.set_seed(1)
        m.set_device(dev)
        m.initialize()
        opt = nn.Adam(1e-4, epsilon=1e-3, weight_decay=1e-3)
        return m, opt, DataParallel(m, opt, dist.group.WORLD, sync_bn=False, nccl_baseline=baseline, num_blocks=num_blocks)

    def bench(dp, m, iters=20, warm=5):
        for g in m.gradients():
            g.normal_()
        for _ in range(warm):
            dp.reduce_and_update()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(iters):
            dp.reduce_and_update()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    m, opt, dp = build(False)
    fused_ms = bench(dp, m)
    kernel_us = {}
    if os.environ.get("AR_PROFILE", "1") == "1":
        # where the time of one call goes: device time per kernel (CUPTI through torch.profiler), rank 0
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(10):
                dp.reduce_and_update()
            torch.cuda.synchronize()
        for ev in prof.key_averages():
            t = getattr(ev, "device_time_total", None) or getattr(ev, "cuda_time_total", 0.0)
            if t and ev.count:
                kernel_us[ev.key[:60]] = round(t / ev.count, 1)
        dist.barrier()
    nbytes = dp.bytes_over_nvlink_per_step()
    total_params = dp.n
    del dp, m, opt
    torch.cuda.empty_cache()
    # CTAs of the fused kernel: the default is one per SM; more resident CTAs keep more NVLink loads in flight
    sweep = {}
    sms = torch.cuda.get_device_properties(local).multi_processor_count
    for mult in (2, 4):
        m, opt, dp = build(False, num_blocks=mult * sms)
        sweep[f"{mult}x_sms"] = round(bench(dp, m), 4)
        del dp, m, opt
        torch.cuda.empty_cache()
    m2, opt2, dp2 = build(True)
    nccl_ms = bench(dp2, m2)
    if rank == 0:
        bw = 770e9      # measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900
        try:
            peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                "MEASURED_PEAKS.json")))
            for k, v in peaks.items():
                if "nvlink" in k.lower() and isinstance(v, (int, float)):
                    bw = float(v) * (1e9 if v < 1e6 else 1.0)
        except (OSError, ValueError):
            pass
        # the two directions overlap, so the floor is max(in, out) = nbytes / bw; the local Adam pass adds
        # 7 * 4 * n / R bytes of HBM traffic (g, w, m, v read; w, m, v written) at ~6.5 TB/s
        floor_ms = nbytes / bw * 1e3
        hbm_ms = 7 * 4 * total_params / world / 6.5e12 * 1e3
        print(json.dumps({"world": world, "params": total_params, "nvlink_bytes_per_dir": nbytes,
                          "fused_ms": round(fused_ms, 4), "fused_kernel_us": kernel_us, "fused_ms_by_ctas": sweep, "nccl_allreduce_plus_update_ms": round(nccl_ms, 4),
                          "nvlink_floor_ms": round(floor_ms, 4), "hbm_floor_ms": round(hbm_ms, 4),
                          "fraction_of_nvlink_roofline": round(floor_ms / fused_ms, 3),
                          "speedup_vs_nccl_path": round(nccl_ms / fused_ms, 2), "nvlink_GBps_assumed": bw / 1e9}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
