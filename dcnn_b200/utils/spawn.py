"""Deadline-bounded multi-process launch: like ``torch.multiprocessing.spawn(join=True)`` but a rank that never returns
(a collective whose peer died, a pre-posted receive that is never matched) is killed -- by PID, only the children this
call started -- after ``timeout_s`` and reported as a ``TimeoutError`` instead of hanging the caller (or a GPU box)."""
from __future__ import annotations

import time

import torch.multiprocessing as mp


def spawn_bounded(fn, args=(), nprocs: int = 1, timeout_s: float = 300.0):
    ctx = mp.spawn(fn, args=args, nprocs=nprocs, join=False)
    deadline = time.monotonic() + timeout_s
    try:
        while not ctx.join(timeout=1.0):          # raises ProcessRaisedException / ProcessExitedException on failure
            if time.monotonic() > deadline:
                raise TimeoutError(f"{nprocs} spawned ranks did not finish within {timeout_s:.0f} s")
    finally:
        for p in ctx.processes:
            if p.is_alive():
                p.kill()
        for p in ctx.processes:
            p.join(5)
