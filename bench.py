#!/usr/bin/env python
"""Headline benchmark: ResNet-50 (Tiny-ImageNet shape, 64x64) training throughput in images/s.

Contract (see the round brief): ``python bench.py --gpus N --steps K --warmup W`` (launched under torchrun for
N>1, one rank per GPU, NCCL for the plumbing).  W untimed warm-up steps, then exactly K timed steps bracketed by
barrier + cuda synchronize, device-timed with CUDA events, max over ranks; rank 0 prints ONE JSON line.
``--impl reference`` runs the unmodified reference build under baseline/_ref for the same metric/config.
Synthetic data of the named shape, random-init weights; compute dtype: fp32 storage, TF32 tensor-core MMA with
FP32 accumulation (the reference is FP32 cuDNN/cuBLAS, which uses TF32 tensor cores for FP32 convolutions)."""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODEL = "ResNet-50-Tiny-ImageNet"
IMG = (3, 64, 64)
CLASSES = 200


class ClockSampler(threading.Thread):
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    def __init__(self, gpu_index: int, uuid=None):
        super().__init__(daemon=True)
        self.gpu_index, self.uuid, self.source = gpu_index, (str(uuid) if uuid else None), "nvidia-smi"
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.proc = None

    def run(self):
        try:
            if self._run_nvml():
                return
        except Exception:      # any NVML problem: fall back to the nvidia-smi loop below
            pass
        self._run_smi()

    def _run_nvml(self) -> bool:
        """In-process NVML polling every 20 ms (nvidia-smi takes a few hundred ms just to start, which is most of a
        short timed region).  Returns False if NVML cannot be used."""
        import pynvml
        pynvml.nvmlInit()
        h = None
        if self.uuid:
            for cand in (self.uuid, "GPU-" + self.uuid):
                try:
                    h = pynvml.nvmlDeviceGetHandleByUUID(cand.encode() if hasattr(cand, "encode") else cand)
                    break
                except Exception:
                    h = None
        if h is None:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)
        self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(pynvml, "nvmlDeviceGetCurrentClocksThrottleReasons")
        bits = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        self.source = "nvml"
        while not self._stop_evt.is_set():
            self.samples.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
            r = int(get_reasons(h))
            for b, n in bits.items():
                if r & b:
                    self.reasons.add(n)
            self._stop_evt.wait(0.02)
        return True

    def _run_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu_index), "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            if self._stop_evt.is_set():
                break
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                self.samples.append(float(parts[0]))
                self.max_mhz = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    self.reasons.add(n)

    def stop(self):
        self._stop_evt.set()
        if self.proc is not None:
            self.proc.terminate()

    def summary(self):
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s),
                "source": self.source}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)     # BASELINE.md's protocol: >= 20 warm-up, >= 100 timed steps
    ap.add_argument("--warmup", type=int, default=20)     # (about 2 s on one B200, 10 s for the reference arm)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=int(os.environ.get("BENCH_BATCH", 256)), help="per-GPU batch")
    ap.add_argument("--parallelism", default=os.environ.get("BENCH_PARALLELISM", "auto"))
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--model", default=os.environ.get("BENCH_MODEL", "resnet50_tiny_imagenet"),
                    help="dcnn_b200.models.MODEL_ZOO key; the headline benchmark is the default (BASELINE.json)")
    ap.add_argument("--device", default="cuda", choices=["cuda", "cpu"],
                    help="cpu: BASELINE config 1 (MNIST CNN on the CPU, --model mnist_cnn --batch 64) measured for real "
                         "on this library's native CPU kernels; the headline benchmark is always cuda")
    ap.add_argument("--microbatches", type=int, default=1)
    return ap.parse_args()


def model_spec(name):
    """(display name, image shape, classes, optimizer constructor arguments as in the reference's example for that model,
    optimizer description) -- the other BASELINE.json configs use the same harness with another --model."""
    if name == "resnet50_tiny_imagenet":
        return MODEL, IMG, CLASSES, dict(lr=1e-4, epsilon=1e-3, weight_decay=1e-3), "Adam(lr 1e-4, eps 1e-3, wd 1e-3)"
    if "tiny_imagenet" in name:      # examples/tiny_imagenet_resnet18.cpp / resnet34.cpp
        return name, (3, 64, 64), 200, dict(lr=1e-4, epsilon=1e-7), "Adam(lr 1e-4, eps 1e-7)"
    if "cifar100" in name:
        return name, (3, 32, 32), 100, dict(lr=1e-3, epsilon=1e-8), "Adam(lr 1e-3, eps 1e-8)"
    if "cifar10" in name:            # examples/cifar10_resnet9.cpp
        return name, (3, 32, 32), 10, dict(lr=1e-3, epsilon=1e-8, weight_decay=5e-4), "Adam(lr 1e-3, eps 1e-8, wd 5e-4)"
    if "mnist" in name:
        return name, (1, 28, 28), 10, dict(lr=1e-2, epsilon=1e-8), "Adam(lr 1e-2, eps 1e-8)"
    raise SystemExit(f"--model {name}: no input shape is associated with this model")


def reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:      # launched with torchrun for N > 1: one line, from rank 0
        return 0
    from baseline import reference_bench
    return reference_bench.run(args)


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm(args)
    if args.parallelism.startswith("pp"):
        # no lazy kernel loading in the middle of a pipeline schedule (see PipelineEngine._warm_kernels)
        os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

    import torch
    import torch.distributed as dist
    from dcnn_b200 import _lib, nn
    from dcnn_b200.models import MODEL_ZOO
    from dcnn_b200.train import TrainStep

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N > 1")
    # BENCH_DEVICE=cpu: plumbing check of THIS SCRIPT on a machine without a GPU (tests/test_bench_contract.py).  Every
    # number it prints is meaningless and marked as such; the benchmark proper always runs on CUDA.
    real_cpu = args.device == "cpu"            # a CPU measurement on purpose (BASELINE config 1), not the dry run
    if real_cpu:
        os.environ.setdefault("DCNN_CPU_BACKEND", "native")
    cpu_mode = real_cpu or os.environ.get("BENCH_DEVICE", "cuda") == "cpu"

    class _CpuEvent:
        def record(self):
            self.t = time.perf_counter()

        def elapsed_time(self, other):
            return (other.t - self.t) * 1000.0

    def new_event():
        return _CpuEvent() if cpu_mode else torch.cuda.Event(enable_timing=True)

    def pin(t):
        return t if cpu_mode else t.pin_memory()

    def device_sync():
        if not cpu_mode:
            torch.cuda.synchronize()

    if cpu_mode:
        dev = torch.device("cpu")
    else:
        torch.cuda.set_device(local_rank)
        dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if cpu_mode:
            dist.init_process_group("gloo")
        elif args.parallelism.startswith("pp"):
            # lazy communicator creation: in eager mode (device_id given) NCCL treats every unbatched isend/irecv of a
            # group as a collective serialised with all other ops of that group
            dist.init_process_group("nccl")
        else:
            dist.init_process_group("nccl", device_id=dev)
        if os.environ.get("DCNN_HANG_DUMP_S"):
            import faulthandler
            faulthandler.dump_traceback_later(float(os.environ["DCNN_HANG_DUMP_S"]), exit=True)

    torch.manual_seed(1234)
    if args.model not in MODEL_ZOO:
        raise SystemExit(f"--model {args.model}: not in MODEL_ZOO ({sorted(MODEL_ZOO)})")
    model_name, img, classes, opt_args, opt_desc = model_spec(args.model)
    headline = args.model == "resnet50_tiny_imagenet"
    model = MODEL_ZOO[args.model]()
    model.set_seed(42)           # identical initial weights on every replica
    opt = nn.Adam(opt_args.pop("lr"), **opt_args)   # headline: examples/tiny_imagenet_resnet50.cpp:90
    loss_fn = nn.LogSoftmaxCrossEntropyLoss()

    # ---- parallelism: single | dpR | ppS | ppSxdpR   (rank = replica * S + stage)
    par = args.parallelism
    if par == "auto":
        par = "single" if world == 1 else f"dp{world}"
    S, R = 1, world
    if par.startswith("pp"):
        S = int(par[2:].split("x")[0])
        R = int(par.split("xdp")[1]) if "xdp" in par else 1
    elif par.startswith("dp"):
        S, R = 1, int(par[2:])
    elif par == "single":
        S, R = 1, 1
    if S * R != world:
        raise SystemExit(f"--parallelism {par} needs {S * R} GPUs, got {world}")
    parallelism = par
    engine = None
    dp = None
    sync_bn = os.environ.get("BENCH_SYNC_BN", "1") == "1"
    if S > 1:
        from dcnn_b200.parallel.pipeline import PipelineEngine, make_pipeline_groups
        from dcnn_b200.partitioner import BalancedPartitioner
        stage, rep, links, dp_g = make_pipeline_groups(S, R)
        mb = args.microbatches if args.microbatches > 1 else 2 * S
        if os.environ.get("BENCH_PARTITION", "flops") == "measured":
            # balance the stages on measured per-layer GPU time of one microbatch (rank 0 measures, everyone uses its
            # numbers) instead of on FLOP counts -- BatchNorm / pooling layers cost time, not FLOPs
            from dcnn_b200.partitioner import measure_layer_costs
            costs = [None]
            if rank == 0:
                probe = MODEL_ZOO[args.model]()
                probe.set_seed(42)
                probe.set_device(dev)
                probe.initialize()
                costs = [measure_layer_costs(probe, [max(1, args.batch * S // mb), *img])]
                del probe
                if not cpu_mode:
                    torch.cuda.empty_cache()
            dist.broadcast_object_list(costs, src=0)
            partitioner = BalancedPartitioner(costs=costs[0])
        else:
            partitioner = BalancedPartitioner(input_shape=[args.batch, *img])
        parts = partitioner.get_partitions(model.get_layers(), S)
        engine = PipelineEngine(model, parts, stage, opt, loss_fn, mb, links, dev,
                                schedule=os.environ.get("SCHEDULE", "semi_async"), dp_group=dp_g if R > 1 else None,
                                sync_bn=sync_bn, seed=42)
        args.microbatches = mb

        class _Step:
            _graph = None

            def __call__(self, x, t):
                return engine.train_step(x, t)

            def _eager(self, x, t):
                return engine.train_step(x, t)
        step = _Step()
    else:
        model.set_device(dev)
        model.initialize()
        model.compute_input_gradient = False
        if world > 1:
            from dcnn_b200.parallel.dp import DataParallel
            dp = DataParallel(model, opt, dist.group.WORLD, sync_bn=sync_bn,
                              nccl_baseline=os.environ.get("BENCH_NCCL_BASELINE", "0") == "1")
        else:
            opt.attach(model.parameters(), model.gradients())
        step = TrainStep(model, opt, loss_fn, use_cuda_graph=not args.no_graph,
                         num_microbatches=args.microbatches, data_parallel=dp)

    B = args.batch * S          # batch fed to one pipeline replica: per-GPU work stays fixed
    g = torch.Generator().manual_seed(1000 + rank // S)
    nbuf = 4
    host_x = [pin(torch.randn(B, *img, generator=g)) for _ in range(nbuf)]
    host_t = [pin(torch.nn.functional.one_hot(torch.randint(0, classes, (B,), generator=g), classes).float()
                  .reshape(B, classes, 1, 1)) for _ in range(nbuf)]
    dev_x = [x.to(dev) for x in host_x]
    dev_t = [t.to(dev) for t in host_t]

    def barrier():
        if world > 1:
            dist.barrier()
        device_sync()

    # ---- kernel-only arm: inputs already resident, no host read-back inside the timed region
    # at least 3 untimed steps whatever --warmup says: the first two run eagerly (lazy init, pointer tables) and the
    # third captures the CUDA graph -- none of that may land in the timed region
    for i in range(max(args.warmup, 3)):
        step(dev_x[i % nbuf], dev_t[i % nbuf])
    barrier()
    sampler = ClockSampler(local_rank if "CUDA_VISIBLE_DEVICES" not in os.environ else 0,
                           None if cpu_mode else getattr(torch.cuda.get_device_properties(dev), "uuid", None))
    if rank == 0 and not cpu_mode:         # no GPU clocks to sample on a CPU run (and the sampler would compete for cores)
        sampler.start()
    _lib.reset_launch_count()
    e0, e1 = new_event(), new_event()
    barrier()
    e0.record()
    for i in range(args.steps):
        loss, correct = step(dev_x[i % nbuf], dev_t[i % nbuf])
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = _lib.launch_count()
    if step._graph is not None:   # replayed graph: count the kernels captured in one step
        launches = step_launches(step, dev_x[0], dev_t[0], _lib) * args.steps
    elif engine is not None and getattr(engine, "_graphs", None) is not None:
        # pipeline graph mode: the per-op graphs' kernels (counted at capture) + what still runs eagerly each step
        launches += engine.graph_launches * args.steps
    final_loss = float(loss) if loss is not None else float('nan')
    stage_stats = engine.stats() if engine is not None else None

    # ---- end-to-end arm: public API, pinned-host -> device copy of every batch + loss read-back every step
    h2d = host_x[0].numel() * 4 + host_t[0].numel() * 4
    d2h = 8
    out_host = [pin(torch.zeros(2, dtype=torch.float32)) for _ in range(2)]
    # The batches come through the library's PrefetchLoader, the streaming front end of the training loops
    # (PREFETCH_DEPTH=k in train_classification_model, dcnn_b200/train.py): a worker thread stages each batch in pinned
    # memory and issues its host-to-device copy
    # on a side stream, so the copy of batch i+1 overlaps the kernels of step i.  One "epoch" of exactly `steps` batches
    # is produced inside the timed region (the loader is started after the first event), nothing is copied ahead of it.
    class _HostBatches:
        """Loader interface over the synthetic pinned host batches: an epoch of `n` batches."""

        def __init__(self, n):
            self.n, self.i = n, 0

        def size(self):
            return self.n * B

        def reset(self):
            self.i = 0

        def get_batch(self, batch_size):
            if self.i >= self.n:
                return None
            k = self.i % nbuf
            self.i += 1
            return host_x[k], host_t[k]

    from dcnn_b200.data import PrefetchLoader
    batches = _HostBatches(0)
    feed = PrefetchLoader(batches, B, depth=2, device=None if cpu_mode else dev)

    def e2e_epoch(n):
        batches.n = n
        feed.reset()
        done, last, landed = 0, (None, None), [None, None]
        while True:
            b = feed.get_next_batch()
            if b is None:
                break
            loss, correct = step(b[0], b[1])
            slot = done & 1
            if loss is not None:                       # every step's loss and correct count are read back ...
                out_host[slot][0].copy_(loss, non_blocking=True)
                if correct is not None:
                    out_host[slot][1].copy_(correct, non_blocking=True)
            if not cpu_mode:
                landed[slot] = new_event()
                landed[slot].record()
                if landed[slot ^ 1] is not None:        # ... and the host waits for the PREVIOUS step's pair once this
                    landed[slot ^ 1].synchronize()      # step is submitted, so the device never waits for the host
            done, last = done + 1, (loss, correct)
        if not cpu_mode:
            torch.cuda.current_stream().synchronize()
        assert done == n, (done, n)
        return last

    e2e_epoch(6)       # warm-up epoch: the loader's pinned staging ring and its side-stream device buffers exist afterwards
    barrier()
    f0, f1 = new_event(), new_event()
    f0.record()
    loss, correct = e2e_epoch(args.steps)
    f1.record()
    barrier()
    feed.close()
    ms_e2e = f0.elapsed_time(f1)
    if rank == 0:
        sampler.stop()
        if sampler.is_alive():
            sampler.join(timeout=2)

    t = torch.tensor([ms, ms_e2e, final_loss if final_loss == final_loss else -1e30, float(launches)], device=dev,
                     dtype=torch.float64)
    bubble = None
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if stage_stats is not None:
            b = torch.tensor([stage_stats["busy_ms"], stage_stats["step_ms"]], device=dev, dtype=torch.float64)
            allb = [torch.zeros_like(b) for _ in range(world)]
            dist.all_gather(allb, b)
            busy = sum(float(v[0]) for v in allb)
            stepm = max(float(v[1]) for v in allb)
            bubble = 1.0 - busy / (world * stepm)
    ms, ms_e2e, final_loss, launches = float(t[0]), float(t[1]), float(t[2]), int(t[3])
    gb = B * R
    value = gb * args.steps / (ms / 1000.0)
    e2e = gb * args.steps / (ms_e2e / 1000.0)
    out = None
    if rank == 0:
        base = None
        try:
            with open(os.path.join(ROOT, "BASELINE.json")) as f:
                pub = json.load(f).get("published", {})
            base = pub.get("images_per_sec") if isinstance(pub, dict) else None
        except (OSError, ValueError):
            pass
        out = {
            "metric": "train images/sec ResNet-50 64x64 (Tiny-ImageNet shape), max over ranks" if headline
            else f"train images/sec {model_name} {img[1]}x{img[2]}, max over ranks",
            "value": round(value, 2), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": (value / base) if base else None,
            "dtype": "fp32 storage, tf32 tensor-core MMA, fp32 accumulate" if not cpu_mode
            else (f"fp32 (CPU, {os.environ.get('DCNN_CPU_BACKEND', 'aten')} kernels, {torch.get_num_threads()} threads)"
                  if real_cpu else "CPU DRY RUN OF THE SCRIPT -- not a measurement"), "data": "synthetic", "impl": "ours",
            # `config` has the same keys (and, for a single device, the same values) on both arms: what is specific to
            # this implementation's run sits beside it in `run`
            "config": {"model": model_name, "global_batch": gb, "per_gpu_batch": B, "seq_len": None,
                       "image": "x".join(map(str, img)), "classes": classes, "parallelism": parallelism,
                       "optimizer": opt_desc, "loss": "LogSoftmaxCrossEntropy",
                       "microbatches": args.microbatches, "pipeline_stages": S, "dp_replicas": R,
                       "l2": "per-step working set (activations, >10 GB at batch 256) is far larger than the 126 MB L2; "
                             "inputs rotate over 4 distinct batches"},
            "run": {"cuda_graph": (step._graph is not None) or
                                  (engine is not None and getattr(engine, "_graphs", None) is not None),
                    "sync_bn": sync_bn and R > 1},
            "pipeline_bubble_fraction": bubble,
            "clocks": sampler.summary(),
            "e2e": {"value": round(e2e, 2), "unit": "images/s", "ms_per_step": round(ms_e2e / args.steps, 4),
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "path": "dcnn_b200.data.PrefetchLoader (pinned staging, H2D on a side stream one batch ahead) -> "
                            "train step -> loss/correct D2H every step, host waits for step i-1's pair after submitting step i"},
            "gpu_launches": int(launches),  # per rank, max over ranks
            "final_loss": final_loss,
        }
    # ---- BASELINE.json's multi-GPU configuration is pipeline x data parallel with a bubble fraction; the headline value
    # above is the data-parallel run (the faster arrangement on one NVSwitch box).  With --gpus N > 1 and the default
    # parallelism, the same N GPUs then run the pipeline arrangement (N = 8: 4 stages x 2 replicas, the named config;
    # N = 4: 4 stages; N = 2: 2 stages) in a fresh process group of child processes, and its line is attached as
    # "pipeline" -- throughput, ms/step and bubble fraction measured by the same protocol.  BENCH_PIPELINE=0 skips it.
    pipe = None
    if world > 1:
        dist.destroy_process_group()
    if world > 1 and args.parallelism == "auto" and not cpu_mode and os.environ.get("BENCH_PIPELINE", "1") == "1" \
            and os.environ.get("BENCH_CHILD") != "1":
        pipe = pipeline_child(args, world, rank)
    if rank == 0:
        if pipe is not None:
            out["pipeline"] = pipe
            if isinstance(pipe, dict) and pipe.get("pipeline_bubble_fraction") is not None:
                out["pipeline_bubble_fraction"] = pipe["pipeline_bubble_fraction"]
        print(json.dumps(out), flush=True)
    if world > 1 and cpu_mode:
        # gloo dry runs: on a loaded host a gloo worker thread now and then drops the last reference to its context during
        # interpreter shutdown and the process dies with "terminate called without an active exception" AFTER the result
        # was printed (seen ~1 in 8 runs with six busy cores).  Everything is torn down and flushed at this point; leave
        # without running the C++ static destructors.  (NCCL runs exit the normal way.)
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def pipeline_child(args, world, rank):
    """Every rank starts one child (same GPU) that joins a new process group on MASTER_PORT + 1 and runs this script
    with --parallelism pp...; rank 0 returns the child's JSON line (or a one-line reason)."""
    par = {2: "pp2", 4: "pp4", 8: "pp4xdp2"}.get(world)
    if par is None:
        return None
    stages = 4 if world >= 4 else 2
    mb = int(os.environ.get("BENCH_PIPELINE_MICROBATCHES", 4 * stages))
    env = dict(os.environ, BENCH_CHILD="1", MASTER_PORT=str(int(os.environ.get("MASTER_PORT", "29500")) + 1),
               DCNN_PIPE_GRAPH=os.environ.get("DCNN_PIPE_GRAPH", "1"),
               BENCH_PARTITION=os.environ.get("BENCH_PARTITION", "measured"))
    for k in list(env):
        if k.startswith("TORCHELASTIC") or k in ("TORCH_NCCL_ASYNC_ERROR_HANDLING",):
            env.pop(k)
    steps = max(5, min(args.steps, 20))
    cmd = [sys.executable, os.path.abspath(__file__), "--gpus", str(world), "--steps", str(steps), "--warmup", "5",
           "--parallelism", par, "--microbatches", str(mb), "--batch", str(args.batch), "--model", args.model]
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=float(os.environ.get("BENCH_PIPELINE_TIMEOUT", 240)))
    except subprocess.TimeoutExpired:
        return {"unavailable": f"{par} child run exceeded its time limit"} if rank == 0 else None
    if rank != 0:
        return None
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not lines:
        tail = (r.stderr.strip().splitlines() or ["no output"])[-1][:200]
        return {"unavailable": f"{par} child run failed: {tail}"}
    d = json.loads(lines[-1])
    return {k: d.get(k) for k in ("value", "unit", "ms_per_step", "steps", "warmup", "pipeline_bubble_fraction", "e2e",
                                  "gpu_launches", "final_loss", "run")} | {
        "parallelism": d["config"]["parallelism"], "microbatches": d["config"]["microbatches"],
        "pipeline_stages": d["config"]["pipeline_stages"], "dp_replicas": d["config"]["dp_replicas"],
        "global_batch": d["config"]["global_batch"], "schedule": os.environ.get("SCHEDULE", "semi_async")}


def step_launches(step, x, t, _lib):
    """Kernels of one training step, counted by running it once eagerly outside the timed region."""
    import torch
    _lib.reset_launch_count()
    step._eager(x, t)
    torch.cuda.synchronize()
    return _lib.launch_count()


if __name__ == "__main__":
    main()
