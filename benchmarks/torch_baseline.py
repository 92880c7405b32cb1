#!/usr/bin/env python
"""The same architectures trained by stock PyTorch (cuDNN / cuBLAS kernels, autograd, torch.optim) -- the role of the
reference's ``torch/torch_resnet9_cifar10.py`` / ``torch_tiny_imagenet_trainer.py`` comparison scripts, and a second
baseline next to the reference arm of bench.py: it answers "what does the vendor library stack reach on this GPU for
this model", which the reference's FP32 NCHW cuDNN path does not.

    python benchmarks/torch_baseline.py [--model resnet50_tiny_imagenet] [--batch 256] [--steps 30] [--warmup 5]
                                        [--tf32 1] [--channels-last 1] [--graph 1] [--device cuda]

The module is produced from the zoo model by ``dcnn_b200.io.to_torch_module`` (identical architecture and initial
weights).  Same protocol as bench.py: synthetic inputs rotating over 4 device-resident batches, Adam with the model's
example hyper-parameters, CUDA-event timing after warm-up.  One JSON line."""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200.io import to_torch_module  # noqa: E402
from dcnn_b200.models import MODEL_ZOO  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50_tiny_imagenet")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--tf32", type=int, default=1, help="allow TF32 tensor-core math in cuDNN / cuBLAS (our kernels use TF32)")
    ap.add_argument("--channels-last", type=int, default=1)
    ap.add_argument("--graph", type=int, default=1, help="capture the step in a CUDA graph (as bench.py does)")
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    args = ap.parse_args()
    dev = torch.device(args.device)
    cuda = dev.type == "cuda"
    if cuda:
        torch.backends.cudnn.allow_tf32 = bool(args.tf32)
        torch.backends.cuda.matmul.allow_tf32 = bool(args.tf32)
        torch.backends.cudnn.benchmark = True
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import model_spec
    name, img, classes, opt_args, opt_desc = model_spec(args.model)
    model = MODEL_ZOO[args.model]()
    model.set_seed(42)
    model.initialize()
    net = to_torch_module(model).to(dev)
    if args.channels_last and cuda:
        net = net.to(memory_format=torch.channels_last)
    net.train()
    lr = opt_args.pop("lr")
    opt = torch.optim.Adam(net.parameters(), lr=lr, eps=opt_args.get("epsilon", 1e-8),
                           weight_decay=opt_args.get("weight_decay", 0.0), capturable=cuda and bool(args.graph))
    g = torch.Generator().manual_seed(1000)
    xs = [torch.randn(args.batch, *img, generator=g).to(dev) for _ in range(4)]
    ts = [torch.randint(0, classes, (args.batch,), generator=g).to(dev) for _ in range(4)]
    if args.channels_last and cuda:
        xs = [x.contiguous(memory_format=torch.channels_last) for x in xs]
    sx, st = xs[0].clone(), ts[0].clone()

    def step_body():
        opt.zero_grad(set_to_none=False)
        loss = torch.nn.functional.cross_entropy(net(sx).flatten(1), st)
        loss.backward()
        opt.step()
        return loss

    graph = None
    if cuda and args.graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                step_body()
        torch.cuda.current_stream().wait_stream(s)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            static_loss = step_body()

    def step(i):
        sx.copy_(xs[i % 4], non_blocking=True)
        st.copy_(ts[i % 4], non_blocking=True)
        if graph is not None:
            graph.replay()
            return static_loss
        return step_body()

    for i in range(max(3, args.warmup)):
        step(i)
    if cuda:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for i in range(args.steps):
        loss = step(i)
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"impl": "pytorch " + torch.__version__ + (" (cuDNN %s)" % torch.backends.cudnn.version() if cuda else " (CPU)"),
                      "metric": f"train images/sec {name}", "value": round(args.batch * args.steps / (ms / 1e3), 1),
                      "unit": "images/s", "ms_per_step": round(ms / args.steps, 3), "steps": args.steps,
                      "warmup": max(3, args.warmup), "batch": args.batch, "tf32": bool(args.tf32),
                      "channels_last": bool(args.channels_last), "cuda_graph": graph is not None,
                      "optimizer": opt_desc, "final_loss": round(float(loss.detach()), 4), "device": str(dev)}))


if __name__ == "__main__":
    main()
