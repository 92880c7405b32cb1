"""How fast can the SMs pull operand tiles through TMA, and what limits it?  (Run on a GPU box; CUDA-event timing, L2
is far smaller than the matrices walked.)  Sweeps: CTAs (1 SM .. all SMs -> per-SM limit vs chip-wide limit), stages
(latency hiding), box shapes of the conv kernels, and a B tile shared by all CTAs (what L2 dedup / multicast exploit)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200._lib import I, LL, P, check, lib  # noqa: E402

dev = "cuda:0"
props = torch.cuda.get_device_properties(0)
sms = props.multi_processor_count
clk = 1.965e9
A = torch.randn(1 << 21, 128, device=dev)       # activations-like: 2M pixels x 128 channels (1 GiB)
B = torch.randn(1 << 14, 128, device=dev)       # weights-like: 8 MiB, L2 resident


def run(a_rows, b_rows, stages, shared, ctas, iters=2000):
    bpi = C.c_longlong(0)

    def go():
        check(lib().dcnn_tma_stream_probe(P(A.data_ptr()), LL(A.shape[0]), LL(A.shape[1]), I(a_rows), P(B.data_ptr()),
                                          LL(B.shape[0]), LL(B.shape[1]), I(b_rows), I(iters), I(stages), I(shared),
                                          I(ctas), C.byref(bpi), P(0)), "tma_stream_probe")
    go()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    go()
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1)
    n = ctas if ctas > 0 else sms
    tot = bpi.value * iters * n
    print(f"A {a_rows:3d} rows  B {b_rows:3d} rows ({'shared' if shared else 'unique'})  stages {stages}  CTAs {n:3d}: "
          f"{ms:7.3f} ms  {tot / ms / 1e9:8.2f} TB/s chip  {tot / n / (ms * 1e-3) / clk:6.1f} B/clk/SM", flush=True)


print(f"{props.name}, {sms} SMs")
for ctas in (1, 8, 37, 74, 148):
    run(128, 128, 4, 1, min(ctas, sms))
for stages in (2, 3, 4, 6):
    run(128, 128, stages, 1, sms)
for (a, b, st) in ((128, 128, 4), (256, 128, 3), (256, 256, 2), (256, 64, 4), (128, 0, 6), (256, 0, 4)):
    for shared in (1, 0):
        if b == 0 and shared == 0:
            continue
        run(a, b, st, shared, sms)
