"""SPMD pipeline engine: the performance path of pipeline parallelism on one 8xB200 box.

Every rank owns one stage (one GPU) and executes a *static schedule table* instead of reacting to message priority
(the reference's effective schedule is emergent, communicator.hpp:100-110):

  * ``sync``        GPipe: all forwards, then all backwards (examples/sync_pipeline_coordinator.cpp:122-180)
  * ``semi_async``  backward of a microbatch starts as soon as its forward leaves the last stage -> 1F1B steady state
                    (coordinator.hpp:273-326), which also bounds live activations to (S - stage) microbatches.

Activations and gradients move GPU->GPU with NCCL p2p (NVLink 5) as dense NHWC buffers.  Every stage-to-stage link has
its own pair of 2-rank communicators (one per direction): NCCL executes the ops of one communicator in issue order,
so a receive that is posted early on a shared communicator would block every later send of that rank behind it and
the 1F1B schedule dead-locks (seen on hardware with one 4-rank group per direction).  With per-link, per-direction
communicators each communicator carries one ordered stream of transfers whose order is identical on both ends.
Receives for the next op are posted before the current op computes, so transfers overlap compute.  The loss lives on the last stage; dL/dinput of the first stage is never computed or shipped.
With ``dp_group`` each stage is replicated and its update is the fused P2P all-reduce+optimizer kernel
(parallel/dp.py); BatchNorm statistics then span the replicas of the stage.

Numerics match the reference: BN statistics per microbatch, weight gradients accumulated over microbatches, each
microbatch gradient scaled by 1/microbatch_size, reported loss = mean over microbatches."""
from __future__ import annotations

import time
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ...nn.loss import Loss
from ...nn.optimizers import Optimizer
from ...nn.sequential import Partition, Sequential
from ...train import split_microbatches
from ...utils import runtime as _runtime


import os as _os

_PIPE_DEBUG = _os.environ.get("DCNN_PIPE_DEBUG", "0") == "1"


def build_schedule(kind: str, stage: int, num_stages: int, num_microbatches: int,
                   warmup_factor: Optional[int] = None) -> List[Tuple[str, int]]:
    """``sync``: all forwards, then all backwards (GPipe).  ``semi_async``: 1F1B -- ``warm`` forwards, then forward /
    backward alternate, then the remaining backwards.  The textbook warm-up, S-1-stage forwards, leaves no slack between
    neighbouring stages: in the steady state every activation / gradient hand-off (NCCL send + receive, 8-33 MB) lands on
    the critical path (measured on 8 GPUs, pp4 x dp2: step 36.9 ms against 29.9 ms for the ideal 1F1B of the same busy
    time).  ``warmup_factor`` = 2 (default, DCNN_PIPE_WARMUP_FACTOR) gives every stage two microbatches of lead over its
    successor instead of one, so a transfer has a whole forward or backward of compute to hide behind; the price is
    activation memory for up to 2(S-1)+1 microbatches on the first stage."""
    m = num_microbatches
    if kind == "sync":
        return [("F", i) for i in range(m)] + [("B", i) for i in range(m)]
    if kind != "semi_async":
        raise ValueError(f"unknown schedule '{kind}' (sync | semi_async)")
    if warmup_factor is None:
        warmup_factor = int(_os.environ.get("DCNN_PIPE_WARMUP_FACTOR", "2") or 2)
    warm = min(max(1, warmup_factor) * (num_stages - 1 - stage), m)
    ops = [("F", i) for i in range(warm)]
    for i in range(m - warm):
        ops.append(("F", warm + i))
        ops.append(("B", i))
    ops += [("B", i) for i in range(m - warm, m)]
    return ops


class PipelineEngine:
    def __init__(self, full_model: Sequential, partitions: Sequence[Partition], stage: int, optimizer: Optimizer,
                 loss_fn: Loss, num_microbatches: int, links: "PipelineLinks", device, schedule: str = "semi_async",
                 dp_group=None, sync_bn: bool = True, seed: Optional[int] = None, nccl_baseline: bool = False):
        self.S, self.stage, self.M = len(partitions), stage, int(num_microbatches)
        self.links = links
        self.device = torch.device(device)
        self.loss_fn, self.optimizer = loss_fn, optimizer
        self.is_first, self.is_last = stage == 0, stage == len(partitions) - 1
        self.partitions = list(partitions)
        self.full_model = full_model
        sub = full_model.split(partitions)[stage]
        if seed is not None:
            sub.set_seed(seed + 1000003 * partitions[stage].start_layer)
        sub.set_device(self.device)
        sub.initialize()
        sub.compute_input_gradient = not self.is_first
        self.model = sub
        self.schedule = build_schedule(schedule, stage, self.S, self.M)
        self.dp = None
        if dp_group is not None and dist.get_world_size(dp_group) > 1:
            from ..dp import DataParallel
            self.dp = DataParallel(sub, optimizer, dp_group, sync_bn=sync_bn, nccl_baseline=nccl_baseline)
        else:
            optimizer.attach(sub.parameters(), sub.gradients())
        self.prev_rank = links.ranks[stage - 1] if not self.is_first else None
        self.next_rank = links.ranks[stage + 1] if not self.is_last else None
        # communicators of this stage's four transfer streams
        self.g_recv_act = links.fwd[stage - 1] if not self.is_first else None   # activations from prev
        self.g_send_act = links.fwd[stage] if not self.is_last else None        # activations to next
        self.g_recv_grad = links.bwd[stage] if not self.is_last else None       # gradients from next
        self.g_send_grad = links.bwd[stage - 1] if not self.is_first else None  # gradients to prev
        self._pending_sends = []
        self.last_stats = {}
        self._kernels_warm = False
        # DCNN_PIPE_GRAPH=1 (CUDA only): after two eager steps every (op, microbatch) of this stage's schedule is
        # captured as its own CUDA graph and replayed from then on; transfers stay outside the graphs.  Opt-in until its
        # first hardware run.
        self.use_graphs = self.device.type == "cuda" and _os.environ.get("DCNN_PIPE_GRAPH", "0") == "1"
        self._graphs = None
        self._steps = 0
        self._warm_links()

    # -- start-up hygiene ------------------------------------------------------------------------------------
    # Two things may block the HOST in the middle of a schedule and, together with a pre-posted receive that is
    # spinning on the GPU, close a cross-rank wait cycle (CUDA documents the first one for lazy module loading):
    #   (a) the first launch of a kernel (lazy module load synchronises the context, i.e. waits for the spinning receive),
    #   (b) the first use of a communicator (connection set-up is a host-side handshake with the peer).
    # Both are therefore done up front, in an order that cannot cycle.
    def _warm_links(self):
        if self.S == 1:
            return
        token = torch.zeros(4, device=self.device)
        for s in range(self.S - 1):          # links in global order: every rank walks the same list
            if self.stage == s:
                dist.send(token, dst=self.links.ranks[s + 1], group=self.links.fwd[s])
                dist.recv(token, src=self.links.ranks[s + 1], group=self.links.bwd[s])
            elif self.stage == s + 1:
                dist.recv(token, src=self.links.ranks[s], group=self.links.fwd[s])
                dist.send(token, dst=self.links.ranks[s], group=self.links.bwd[s])
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def _warm_kernels(self, in_shape, targets_shape):
        """One local forward+backward(+loss) on zeros so that every kernel of this stage is loaded and the caching
        allocator has its blocks before any transfer is posted.  Model state is restored afterwards."""
        if self._kernels_warm:
            return
        self._kernels_warm = True
        if self.device.type != "cuda":
            return
        bufs = {k: v.clone() for k, v in self.model.buffers().items()}
        scratch_mb = self.M + 7
        x = torch.zeros(in_shape, device=self.device)
        y = self.model.forward(x, scratch_mb)
        if self.is_last and targets_shape is not None:
            t = torch.zeros(targets_shape, device=self.device)
            t.view(targets_shape[0], -1)[:, 0] = 1
            _, g, _ = self.loss_fn.forward_backward(y, t)
        else:
            g = torch.zeros_like(y)
        self.model.backward(g, scratch_mb)
        for gr in self.model.gradients():
            gr.zero_()
        for k, v in self.model.buffers().items():
            v.copy_(bufs[k])
        self.model.clear_cache()
        torch.cuda.synchronize(self.device)

    # -- shapes ------------------------------------------------------------------------------------------
    def _boundary_shapes(self, mb_input_shape):
        """Input shape of this stage and of the next stage for one microbatch."""
        s = list(mb_input_shape)
        shapes = [s]
        for part in self.partitions:
            for l in self.full_model.layers[part.start_layer:part.end_layer]:
                s = l.compute_output_shape(s)
            shapes.append(list(s))
        return shapes[self.stage], shapes[self.stage + 1]

    # -- wire --------------------------------------------------------------------------------------------
    def _wire(self, t):
        if t.dim() == 4 and not t.is_contiguous() and t.is_contiguous(memory_format=torch.channels_last):
            return t.permute(0, 2, 3, 1)
        return t.contiguous()

    def _alloc(self, shape):
        if self.device.type == "cuda" and len(shape) == 4:
            n, c, h, w = shape
            buf = torch.empty((n, h, w, c), device=self.device, dtype=torch.float32)
            return buf, buf.permute(0, 3, 1, 2)
        buf = torch.empty(shape, device=self.device, dtype=torch.float32)
        return buf, buf

    def _isend(self, t, dst, group):
        buf = self._wire(t)
        if self.device.type != "cuda" and buf.dim() == 4 and tuple(buf.shape) != tuple(t.shape):
            buf = t.contiguous()
        self._pending_sends.append((dist.isend(buf, dst=dst, group=group), buf))

    def _irecv(self, shape, src, group):
        buf, view = self._alloc(shape)
        return dist.irecv(buf, src=src, group=group), view

    # -- one optimisation step -------------------------------------------------------------------------------
    def train_step(self, x: Optional[torch.Tensor] = None, targets: Optional[torch.Tensor] = None):
        """x: this replica's batch (needed on the first stage), targets: needed on the last stage.
        Returns (mean loss tensor | None, correct-count tensor | None) -- non-None on the last stage."""
        M = self.M
        _runtime.kick()                   # feeds any running StepWatchdog (STEP_TIMEOUT_S)
        timing = self.device.type == "cuda"
        xs = split_microbatches(x.to(self.device, non_blocking=True), M) if self.is_first else None
        ts = split_microbatches(targets.to(self.device, non_blocking=True), M) if self.is_last else None
        if self.is_first:
            mb_shapes = [list(t.shape) for t in xs]
            meta = torch.tensor([s[0] for s in mb_shapes] + list(mb_shapes[0][1:]), dtype=torch.int64)
        else:
            meta = None
        mb_batch, sample_shape = self._microbatch_layout(meta, M)
        in_shapes, out_shapes = [], []
        for b in mb_batch:
            i, o = self._boundary_shapes([b] + sample_shape)
            in_shapes.append(i)
            out_shapes.append(o)

        self._warm_kernels(in_shapes[0], list(ts[0].shape) if self.is_last else None)
        sched = self.schedule
        if self.use_graphs:
            self._steps += 1
            if self._graphs is None and self._steps > 2:       # every rank reaches this on the same step
                self._capture_graphs(in_shapes, out_shapes, [list(t.shape) for t in ts] if self.is_last else None)
            if self._graphs is not None:
                return self._train_step_graphs(xs, ts, in_shapes, out_shapes)
        recv_slot = {}

        def post_recv(k):
            if k >= len(sched):
                return
            op, mb = sched[k]
            if op == "F" and not self.is_first:
                recv_slot[k] = self._irecv(in_shapes[mb], self.prev_rank, self.g_recv_act)
            elif op == "B" and not self.is_last:
                recv_slot[k] = self._irecv(out_shapes[mb], self.next_rank, self.g_recv_grad)

        loss_acc = correct_acc = None
        grads = {}
        busy = []
        t_begin = time.perf_counter()
        if timing:
            ev_begin = torch.cuda.Event(enable_timing=True)
            ev_begin.record()
        post_recv(0)
        debug = _PIPE_DEBUG
        for k, (op, mb) in enumerate(sched):
            if debug:
                print(f"[pipe s{self.stage}] op {k}: {op}{mb}", flush=True)
            post_recv(k + 1)       # overlap the next transfer with this op's compute
            if k in recv_slot:
                work, view = recv_slot.pop(k)
                work.wait()
            if timing:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            else:
                c0 = time.perf_counter()
            if op == "F":
                inp = xs[mb] if self.is_first else view
                y = self.model.forward(inp, mb)
                if self.is_last:
                    loss, g, correct = self.loss_fn.forward_backward(y, ts[mb])
                    grads[mb] = g
                    loss_acc = loss if loss_acc is None else loss_acc + loss
                    if correct is not None:
                        correct_acc = correct if correct_acc is None else correct_acc + correct
                else:
                    self._isend(y, self.next_rank, self.g_send_act)
            else:
                g = grads.pop(mb) if self.is_last else view
                if self.dp is not None:      # last backward of the step: bucketed DP may start reducing early
                    self.dp.final_backward = all(o != "B" for o, _ in sched[k + 1:])
                dx = self.model.backward(g, mb)
                if not self.is_first:
                    self._isend(dx, self.prev_rank, self.g_send_grad)
            if timing:
                e1.record()
                busy.append((e0, e1))
            else:
                busy.append(time.perf_counter() - c0)
        if self.dp is not None:
            self.dp.reduce_and_update(self.optimizer)
        else:
            self.optimizer.update(zero_grads=True)
        for w, _ in self._pending_sends:
            w.wait()
        self._pending_sends = []
        if timing:
            ev_end = torch.cuda.Event(enable_timing=True)
            ev_end.record()
            self._events = (ev_begin, ev_end, busy)
        else:
            self.last_stats = {"busy_ms": sum(busy) * 1000, "step_ms": (time.perf_counter() - t_begin) * 1000}
        if self.is_last and loss_acc is not None and M > 1:
            loss_acc = loss_acc / M
        return loss_acc, correct_acc

    # -- checkpoint of a pipelined run ---------------------------------------------------------------------------
    def gather_full_model(self, dst_stage: int = 0) -> Optional[Sequential]:
        """Collective over the stages of this replica: returns, on stage ``dst_stage``, a CPU :class:`Sequential` with
        the whole model's structure and the parameters / BatchNorm statistics currently held by the stages (None on the
        other stages).  ``model.save_to_file(path)`` on it writes the reference-compatible ``.json`` + ``.bin`` -- the
        reference itself can never gather weights trained through its pipeline."""
        mine = {"stage": self.stage,
                "params": [p.detach().to("cpu").contiguous() for p in self.model.parameters()],
                "buffers": [v.detach().to("cpu") for v in self.model.buffers().values()]}
        if self.S == 1:
            shards = [mine]
        else:
            shards = [None] * self.S
            dist.all_gather_object(shards, mine, group=self.links.ctl)
        if self.stage != dst_stage:
            return None
        full = Sequential.load_from_config(self.full_model.get_config())
        full.initialize()
        shards.sort(key=lambda d: d["stage"])
        full.load_parameters([p for d in shards for p in d["params"]])
        bufs = [b for d in shards for b in d["buffers"]]
        own = list(full.buffers().values())
        if len(own) != len(bufs):
            raise RuntimeError("buffer count mismatch while gathering the pipelined model")
        for dst, src in zip(own, bufs):
            dst.copy_(src)
        return full

    # -- forward-only pass (validation) ------------------------------------------------------------------------
    def eval_step(self, x: Optional[torch.Tensor] = None, targets: Optional[torch.Tensor] = None):
        """Forward the batch through the pipeline in evaluation mode (BatchNorm running statistics, no dropout), loss and
        correct count on the last stage.  Returns (mean loss | None, correct count | None) like ``train_step``."""
        M = self.M
        xs = split_microbatches(x.to(self.device, non_blocking=True), M) if self.is_first else None
        ts = split_microbatches(targets.to(self.device, non_blocking=True), M) if self.is_last else None
        meta = None
        if self.is_first:
            shapes = [list(t.shape) for t in xs]
            meta = torch.tensor([s[0] for s in shapes] + list(shapes[0][1:]), dtype=torch.int64)
        mb_batch, sample_shape = self._microbatch_layout(meta, M)
        in_shapes = [self._boundary_shapes([b] + sample_shape)[0] for b in mb_batch]
        self._warm_kernels(in_shapes[0], list(ts[0].shape) if self.is_last else None)
        was_training = self.model.is_training()
        self.model.set_training(False)
        loss_acc = correct_acc = None
        try:
            nxt = self._irecv(in_shapes[0], self.prev_rank, self.g_recv_act) if not self.is_first else None
            for mb in range(len(mb_batch)):
                cur = nxt
                if not self.is_first and mb + 1 < len(mb_batch):
                    nxt = self._irecv(in_shapes[mb + 1], self.prev_rank, self.g_recv_act)
                if cur is not None:
                    cur[0].wait()
                y = self.model.forward(xs[mb] if self.is_first else cur[1], mb)
                if self.is_last:
                    loss, _, correct = self.loss_fn.forward_backward(y, ts[mb], want_grad=False)
                    loss_acc = loss if loss_acc is None else loss_acc + loss
                    if correct is not None:
                        correct_acc = correct if correct_acc is None else correct_acc + correct
                else:
                    self._isend(y, self.next_rank, self.g_send_act)
            for w, _ in self._pending_sends:
                w.wait()
            self._pending_sends = []
        finally:
            self.model.set_training(was_training)
            self.model.clear_cache()
        if self.is_last and loss_acc is not None:
            loss_acc = loss_acc / len(mb_batch)
        return loss_acc, correct_acc

    # -- CUDA-graph mode ---------------------------------------------------------------------------------------
    def _capture_graphs(self, in_shapes, out_shapes, target_shapes):
        """One CUDA graph per schedule entry, captured in schedule order into one memory pool (graphs that share a pool
        must be replayed in capture order, which is the schedule order).  Inputs live in static buffers: the activation
        / gradient a transfer delivers, the first stage's microbatch, the last stage's targets.  Nothing executes and
        nothing is communicated during capture; the calling step is then run through the replay path."""
        torch.cuda.synchronize(self.device)
        M = self.M
        st = {"in": [self._alloc(in_shapes[m]) for m in range(M)],
              "gin": [self._alloc(out_shapes[m]) if not self.is_last else None for m in range(M)],
              "t": [torch.zeros(target_shapes[m], device=self.device) for m in range(M)] if self.is_last else None,
              "y": {}, "dx": {}, "g": {}, "loss": {}, "correct": {}, "graph": {}}
        if self.is_last:
            for t in st["t"]:
                t.view(t.shape[0], -1)[:, 0] = 1          # a valid one-hot target while capturing
        pool = torch.cuda.graph_pool_handle()
        from ... import _lib
        launches_before = _lib.launch_count()
        for (op, mb) in self.schedule:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, pool=pool):
                if op == "F":
                    y = self.model.forward(st["in"][mb][1], mb)
                    st["y"][mb] = y
                    if self.is_last:
                        loss, gr, correct = self.loss_fn.forward_backward(y, st["t"][mb])
                        st["loss"][mb], st["g"][mb], st["correct"][mb] = loss, gr, correct
                else:
                    gin = st["g"][mb] if self.is_last else st["gin"][mb][1]
                    st["dx"][mb] = self.model.backward(gin, mb)
            st["graph"][(op, mb)] = g
        torch.cuda.synchronize(self.device)
        self.graph_launches = _lib.launch_count() - launches_before    # kernels one replayed step runs from the graphs
        self._graphs = st

    def _train_step_graphs(self, xs, ts, in_shapes, out_shapes):
        st, sched, M = self._graphs, self.schedule, self.M
        recv_slot = {}

        def post_recv(k):
            if k >= len(sched):
                return
            op, mb = sched[k]
            if op == "F" and not self.is_first:
                recv_slot[k] = dist.irecv(st["in"][mb][0], src=self.prev_rank, group=self.g_recv_act)
            elif op == "B" and not self.is_last:
                recv_slot[k] = dist.irecv(st["gin"][mb][0], src=self.next_rank, group=self.g_recv_grad)

        busy = []
        ev_begin = torch.cuda.Event(enable_timing=True)
        ev_begin.record()
        post_recv(0)
        for k, (op, mb) in enumerate(sched):
            post_recv(k + 1)
            if k in recv_slot:
                recv_slot.pop(k).wait()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if op == "F":
                if self.is_first:
                    st["in"][mb][1].copy_(xs[mb], non_blocking=True)
                if self.is_last:
                    st["t"][mb].copy_(ts[mb], non_blocking=True)
                st["graph"][(op, mb)].replay()
                if not self.is_last:
                    self._isend(st["y"][mb], self.next_rank, self.g_send_act)
            else:
                st["graph"][(op, mb)].replay()
                if not self.is_first:
                    self._isend(st["dx"][mb], self.prev_rank, self.g_send_grad)
            e1.record()
            busy.append((e0, e1))
        if self.dp is not None:
            self.dp.reduce_and_update(self.optimizer)
        else:
            self.optimizer.update(zero_grads=True)
        for w, _ in self._pending_sends:       # the static output buffers are rewritten by the next step's replays
            w.wait()
        self._pending_sends = []
        ev_end = torch.cuda.Event(enable_timing=True)
        ev_end.record()
        self._events = (ev_begin, ev_end, busy)
        loss_acc = correct_acc = None
        if self.is_last:
            loss_acc = torch.stack([st["loss"][m] for m in range(M)]).sum() / M
            if all(st["correct"][m] is not None for m in range(M)):
                correct_acc = torch.stack([st["correct"][m] for m in range(M)]).sum()
        return loss_acc, correct_acc

    def stats(self):
        """(busy_ms, step_ms) of the last step on this stage; synchronises."""
        if self.device.type == "cuda" and getattr(self, "_events", None):
            torch.cuda.synchronize(self.device)
            b, e, busy = self._events
            self.last_stats = {"busy_ms": sum(a.elapsed_time(z) for a, z in busy), "step_ms": b.elapsed_time(e)}
        return self.last_stats

    def _microbatch_layout(self, meta, M):
        """Every stage needs the microbatch sizes and the sample shape; stage 0 knows them, the rest learn them once
        (cached) through the forward group."""
        key = "_layout"
        if self.is_first:
            vals = meta.tolist()
            layout = (vals[:M], vals[M:])
            if getattr(self, key, None) != layout:
                setattr(self, key, layout)
                obj = [layout]
            else:
                obj = [None]
        else:
            obj = [None]
        if self.S > 1 and not getattr(self, "_layout_synced", False):
            # one-time exchange (batch geometry is static afterwards; call reset_layout() if it changes)
            dist.broadcast_object_list(obj, src=self.links.ranks[0], group=self.links.ctl)
            self._layout_synced = True
        if obj[0] is not None:
            setattr(self, key, (list(obj[0][0]), list(obj[0][1])))
        layout = getattr(self, key)
        return list(layout[0]), list(layout[1])


class PipelineLinks:
    """Communicators of one pipeline replica: ``fwd[s]`` / ``bwd[s]`` are 2-rank groups for the link between stage s
    and s+1 (one per direction), ``ctl`` spans all stages (set-up only), ``ranks[s]`` is stage s's global rank."""

    def __init__(self, ranks, fwd, bwd, ctl):
        self.ranks, self.fwd, self.bwd, self.ctl = list(ranks), fwd, bwd, ctl

    def reset(self):
        pass


def make_pipeline_groups(num_stages: int, dp_replicas: int, backend: Optional[str] = None):
    """Rank layout: rank = replica * S + stage.  Returns (stage, replica, links, dp_group).
    Every rank must call this (group creation is collective over the whole job)."""
    rank, world = dist.get_rank(), dist.get_world_size()
    if num_stages * dp_replicas != world:
        raise ValueError(f"stages ({num_stages}) x replicas ({dp_replicas}) != world size ({world})")
    my_stage, my_rep = rank % num_stages, rank // num_stages
    links = dpg = None
    for r in range(dp_replicas):
        ranks = [r * num_stages + s for s in range(num_stages)]
        ctl = dist.new_group(ranks=ranks, backend=backend)
        fwd, bwd = {}, {}
        for s in range(num_stages - 1):
            fwd[s] = dist.new_group(ranks=[ranks[s], ranks[s + 1]], backend=backend)
            bwd[s] = dist.new_group(ranks=[ranks[s], ranks[s + 1]], backend=backend)
        if r == my_rep:
            links = PipelineLinks(ranks, fwd, bwd, ctl)
    for s in range(num_stages):
        ranks = [r * num_stages + s for r in range(dp_replicas)]
        g = dist.new_group(ranks=ranks, backend=backend)
        if s == my_stage:
            dpg = g
    return my_stage, my_rep, links, dpg
