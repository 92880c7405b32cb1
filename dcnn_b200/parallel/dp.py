"""Stage-replicated data parallelism over NVLink peer memory (new capability: the reference has no DP at all,
SURVEY §2.3; BASELINE.json makes it the communication hot path).

CUDA path (the product): every replica's parameters, gradients and BatchNorm moment exchange areas live in
*symmetric* buffers (same layout on every rank, cudaMalloc + CUDA IPC so each rank can dereference all of them).
``reduce_and_update`` is ONE hand-written kernel per step (csrc/comm.cu): reduce-scatter by P2P loads of every
replica's gradient slice, 1/R scale, SGD/momentum/Adam/AdamW update of the slice this rank owns (optimizer state is
sharded), all-gather by P2P stores of the new parameters into every replica, then gradient zeroing.  No NCCL call and
no separate elementwise kernel runs on this path.  Sync-BatchNorm merges per-channel moments inside the BN finalize
kernels through the same symmetric memory + epoch flags (csrc/norm.cu).

``nccl_baseline=True`` keeps the conventional NCCL all-reduce + separate optimizer launch as the A/B baseline and
correctness oracle.  CPU tensors (gloo) take the same baseline route so the logic is testable without GPUs."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..nn.layers import BatchNormLayer
from ..nn.optimizers import SGD, Adam, Optimizer
from ..nn.sequential import Sequential


def _all_layers(model):
    from ..nn.blocks import ResidualBlock
    out = []
    for l in model.layers:
        if isinstance(l, ResidualBlock):
            out.extend(l.main_path + l.shortcut_path)
        else:
            out.append(l)
    return out


class _RawCuda:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


class SymmetricBuffer:
    """``nbytes`` of device memory with the same layout on every rank of ``group``; ``peer_ptrs[r]`` is rank r's
    buffer as mapped into this process (CUDA IPC, NVLink peer access)."""

    def __init__(self, nbytes: int, group, device: torch.device):
        from .._lib import check, lib
        self.nbytes = int((nbytes + 255) // 256 * 256)
        self.group, self.device = group, device
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        l = lib()
        p = C.c_void_p()
        check(l.dcnn_dev_malloc(C.byref(p), C.c_longlong(self.nbytes)), "dcnn_dev_malloc")
        self.ptr = p.value
        handle = (C.c_ubyte * 64)()
        check(l.dcnn_ipc_get_handle(C.c_void_p(self.ptr), handle), "dcnn_ipc_get_handle")
        handles = [None] * self.world
        dist.all_gather_object(handles, (bytes(handle), torch.cuda.current_device()), group=group)
        self.peer_ptrs = []
        for r, (h, _dev) in enumerate(handles):
            if r == self.rank:
                self.peer_ptrs.append(self.ptr)
            else:
                q = C.c_void_p()
                hb = (C.c_ubyte * 64).from_buffer_copy(h)
                check(l.dcnn_ipc_open(hb, C.byref(q)), "dcnn_ipc_open")
                self.peer_ptrs.append(q.value)
        self._bytes = torch.as_tensor(_RawCuda(self.ptr, self.nbytes), device=device)

    def view(self, dtype, offset_bytes: int, numel: int) -> torch.Tensor:
        size = torch.empty((), dtype=dtype).element_size()
        return self._bytes[offset_bytes:offset_bytes + numel * size].view(dtype)

    def peer_array(self, offset_bytes: int = 0):
        return (C.c_void_p * self.world)(*[p + offset_bytes for p in self.peer_ptrs])

    def close(self):
        """Collective: every rank closes its mappings of the peers' buffers, then frees its own.  Tensors viewing the
        buffer must not be used afterwards."""
        from .._lib import check, lib
        if self.ptr is None:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        self._bytes = None
        for r, p in enumerate(self.peer_ptrs):
            if r != self.rank:
                check(lib().dcnn_ipc_close(C.c_void_p(p)), "dcnn_ipc_close")
        dist.barrier(group=self.group)            # nobody frees memory that a peer still has mapped
        check(lib().dcnn_dev_free(C.c_void_p(self.ptr)), "dcnn_dev_free")
        self.ptr, self.peer_ptrs = None, []


def send_fd_to_ranks(tag: str, ranks, fd: int, timeout_s: float = 60.0):
    """Hand a file descriptor to other local processes: each receiver listens on the abstract-namespace UNIX socket
    ``\\0<tag>-<rank>`` (see :func:`recv_fd`); the descriptor travels as SCM_RIGHTS ancillary data."""
    import socket
    import time
    for r in ranks:
        addr = "\0" + f"{tag}-{r}"
        deadline = time.monotonic() + timeout_s
        while True:
            s_ = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            try:
                s_.connect(addr)
                break
            except OSError:
                s_.close()
                if time.monotonic() > deadline:
                    raise
                time.sleep(0.02)
        try:
            socket.send_fds(s_, [b"fd"], [fd])
        finally:
            s_.close()


def recv_fd(tag: str, rank: int, timeout_s: float = 60.0) -> int:
    import socket
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    try:
        srv.bind("\0" + f"{tag}-{rank}")
        srv.listen(1)
        srv.settimeout(timeout_s)
        conn, _ = srv.accept()
        try:
            _msg, fds, _flags, _addr = socket.recv_fds(conn, 16, 1)
        finally:
            conn.close()
    finally:
        srv.close()
    if not fds:
        raise RuntimeError("no file descriptor received")
    return fds[0]


class NvlsBuffer:
    """Flat device buffer whose memory is bound to an NVSwitch MULTICAST object spanning the replicas of ``group``
    (csrc/nvls.cu): ``ptr`` is this replica's ordinary (unicast) mapping, ``mc_ptr`` the multicast mapping on which
    ``multimem.ld_reduce`` sums all replicas' copies inside the switch and ``multimem.st`` writes all of them.

    Set-up is collective: rank 0 creates the object and exports a POSIX file descriptor, which reaches the other ranks
    over abstract UNIX sockets (SCM_RIGHTS); every rank adds its device, then (after a barrier) allocates, binds and
    maps.  Opt-in, not yet run on hardware (DCNN_DP_NVLS=1)."""

    def __init__(self, nbytes: int, group, device: torch.device):
        import time
        from .._lib import check, lib
        l = lib()
        l.dcnn_nvls_padded_size.restype = C.c_longlong
        self.group, self.device = group, device
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        dev = device.index if device.index is not None else torch.cuda.current_device()
        self.nbytes = int(l.dcnn_nvls_padded_size(C.c_int(self.world), C.c_longlong(int(nbytes)), C.c_int(dev)))
        if self.nbytes <= 0:
            raise RuntimeError("dcnn_nvls_padded_size failed: " + l.dcnn_last_error().decode())
        handle = C.c_ulonglong(0)
        tag = [None]
        if self.rank == 0:
            import os
            tag[0] = f"dcnn-nvls-{os.getpid()}-{time.monotonic_ns()}"
        dist.broadcast_object_list(tag, src=dist.get_global_rank(group, 0), group=group)
        if self.rank == 0:
            fd = C.c_int(-1)
            check(l.dcnn_nvls_create(C.c_int(self.world), C.c_longlong(self.nbytes), C.byref(handle), C.byref(fd)),
                  "dcnn_nvls_create")
            import os
            send_fd_to_ranks(tag[0], range(1, self.world), fd.value)
            os.close(fd.value)
        else:
            check(l.dcnn_nvls_import(C.c_int(recv_fd(tag[0], self.rank)), C.byref(handle)), "dcnn_nvls_import")
        self.handle = handle.value
        check(l.dcnn_nvls_add_device(C.c_ulonglong(self.handle), C.c_int(dev)), "dcnn_nvls_add_device")
        dist.barrier(group=group)                      # every device is part of the object before anyone binds
        uc, mc, mem = C.c_void_p(), C.c_void_p(), C.c_ulonglong(0)
        check(l.dcnn_nvls_bind(C.c_ulonglong(self.handle), C.c_int(dev), C.c_longlong(self.nbytes), C.byref(uc),
                               C.byref(mc), C.byref(mem)), "dcnn_nvls_bind")
        self.ptr, self.mc_ptr, self.mem_handle = uc.value, mc.value, mem.value
        self._bytes = torch.as_tensor(_RawCuda(self.ptr, self.nbytes), device=device)
        torch.cuda.synchronize(device)
        dist.barrier(group=group)                      # all bindings exist before the first multimem access

    def view(self, dtype, offset_bytes: int, numel: int) -> torch.Tensor:
        size = torch.empty((), dtype=dtype).element_size()
        return self._bytes[offset_bytes:offset_bytes + numel * size].view(dtype)

    def close(self):
        """Collective: unmap both views, release the allocation and the multicast object.  Tensors viewing the buffer
        must not be used afterwards."""
        from .._lib import check, lib
        if self.ptr is None:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        self._bytes = None
        check(lib().dcnn_nvls_release(C.c_ulonglong(self.handle), C.c_ulonglong(self.mem_handle), C.c_void_p(self.ptr),
                                      C.c_void_p(self.mc_ptr), C.c_longlong(self.nbytes)), "dcnn_nvls_release")
        self.ptr = self.mc_ptr = None


class SyncBN:
    """Cross-replica BatchNorm statistics for one BN layer (forward moments and backward sums).

    CUDA: ONE launch per direction -- the finalize kernel pushes its per-channel values into every replica's exchange
    area as flag-in-data words (one NVLink store latency, no fence) and merges all sources from local memory
    (csrc/norm.cu, ``bn_sync_exchange``); the call counters live on the device, so there is no host-side bookkeeping
    to keep in step with CUDA-graph replays."""

    def __init__(self, dp: "DataParallel", channels: int, offset_bytes: int):
        self.dp, self.C = dp, channels
        self.off = offset_bytes          # [forward area][backward area], each area_bytes(world, C)
        if dp.device.type == "cuda":
            a = self.area_bytes(dp.world, channels)
            self.areas_f = dp.bn_buf.peer_array(offset_bytes)
            self.areas_b = dp.bn_buf.peer_array(offset_bytes + a)

    @staticmethod
    def area_bytes(world: int, channels: int) -> int:
        g = (channels + 7) // 8           # must equal dcnn_bn_sync_area_bytes (checked in DataParallel._setup_sync_bn)
        b = 2 * world * 6 * channels * 8 + g * 8
        return (b + 255) // 256 * 256

    @classmethod
    def bytes_needed(cls, world: int, channels: int) -> int:
        return 2 * cls.area_bytes(world, channels)

    def finalize_forward(self, stats, eps, momentum, rm, rv, gamma, beta, mean, invstd, scale, shift):
        from .._lib import F, I, check, lib
        from ..ops.cuda import _p, _stream
        dp = self.dp
        check(lib().dcnn_bn_finalize_sync1(_p(stats.partials), I(stats.tiles), I(self.C), F(eps), F(momentum), _p(rm),
                                           _p(rv), _p(gamma), _p(beta), _p(mean), _p(invstd), _p(scale), _p(shift),
                                           self.areas_f, I(dp.rank), I(dp.world), _stream()), "bn_finalize_sync1")

    def finalize_backward(self, partials, tiles, Cc, M, mean, invstd, gamma, dgamma, dbeta, k1, k2, k3):
        from .._lib import D, I, check, lib
        from ..ops.cuda import _p, _stream
        dp = self.dp
        check(lib().dcnn_bn_bwd_finalize_sync1(_p(partials), I(tiles), I(Cc), D(float(M)), _p(mean), _p(invstd),
                                               _p(gamma), _p(dgamma), _p(dbeta), _p(k1), _p(k2), _p(k3), self.areas_b,
                                               I(dp.rank), I(dp.world), _stream()), "bn_bwd_finalize_sync1")

    # ---- CPU / baseline: the same merges with torch.distributed collectives
    def merge_cpu(self, mean, var, n):
        g = self.dp.group
        cnt = torch.tensor([float(n)], dtype=torch.float64, device=mean.device)
        s = mean.double() * n
        q = (var.double() + mean.double() ** 2) * n
        for t in (cnt, s, q):
            dist.all_reduce(t, group=g)
        N = float(cnt)
        m = s / N
        v = (q / N - m * m).clamp_min(0)
        return m.to(mean.dtype), v.to(var.dtype), int(N)

    def merge_bwd_cpu(self, sum_dy, sum_dy_xhat, m):
        g = self.dp.group
        a, b = sum_dy.clone(), sum_dy_xhat.clone()
        dist.all_reduce(a, group=g)
        dist.all_reduce(b, group=g)
        return a, b, m * self.dp.world


def plan_buckets(unit_params, numels, offsets, n, k, rank, world):
    """Cut the flat gradient buffer ``[0, n)`` into at most ``k`` contiguous buckets at fusion-unit boundaries, balanced
    by parameter count.  ``unit_params[u] = (lo, hi)`` is unit u's parameter-index range, ``offsets[i]`` the flat offset
    of parameter i (multiples of 4), ``numels[i]`` its size.  Returns one dict per bucket, in unit order:
    ``first_unit`` (the unit whose backward completes the bucket), ``a`` / ``n`` (flat start / length, multiple of 4),
    ``sb`` / ``se`` (this rank's shard inside the bucket, bucket-relative, multiples of 4) and ``state_off`` (offset of
    that shard in this rank's optimizer-state arrays)."""
    total = sum(numels)
    target, groups, cur, acc = total / max(1, k), [], [], 0
    for ui, (lo, hi) in enumerate(unit_params):
        cur.append(ui)
        acc += sum(numels[lo:hi])
        if acc >= target * (len(groups) + 1) and len(groups) < k - 1 and hi < len(offsets):
            groups.append(cur)
            cur = []
    if cur:
        groups.append(cur)
    # a group without parameters (pooling, flatten ...) joins its predecessor
    merged = []
    for g in groups:
        if merged and not any(unit_params[u][1] > unit_params[u][0] for u in g):
            merged[-1].extend(g)
        else:
            merged.append(list(g))
    if merged and not any(unit_params[u][1] > unit_params[u][0] for u in merged[0]) and len(merged) > 1:
        merged[1] = merged[0] + merged[1]
        merged = merged[1:]
    buckets, state_off = [], 0
    for gi, g in enumerate(merged):
        first_param = next((unit_params[u][0] for u in g if unit_params[u][1] > unit_params[u][0]), None)
        a = 0 if gi == 0 else offsets[first_param]
        if gi + 1 < len(merged):
            nxt = next(unit_params[u][0] for u in merged[gi + 1] if unit_params[u][1] > unit_params[u][0])
            b = offsets[nxt]
        else:
            b = n
        length = b - a
        if length <= 0 or length % 4:
            raise ValueError(f"bad bucket [{a}, {b})")
        q = length // 4
        sb, se = (q * rank // world) * 4, (q * (rank + 1) // world) * 4
        buckets.append({"first_unit": g[0], "a": a, "n": length, "sb": sb, "se": se, "state_off": state_off})
        state_off += se - sb
    return buckets


def state_layout(n, world, rank, buckets=None):
    """Slices of the flat parameter space whose optimizer state rank ``rank`` keeps, as ``(flat_begin, flat_end, offset
    in the rank's state arrays)``: one contiguous shard, or one shard per bucket (``plan_buckets``)."""
    if buckets is None:
        shard = n // world
        return [(rank * shard, (rank + 1) * shard, 0)]
    out, off = [], 0
    for bk in buckets:
        q = bk["n"] // 4
        sb, se = (q * rank // world) * 4, (q * (rank + 1) // world) * 4
        out.append((bk["a"] + sb, bk["a"] + se, off))
        off += se - sb
    return out


class DataParallel:
    def __init__(self, model: Sequential, optimizer: Optimizer, group=None, sync_bn: bool = True,
                 nccl_baseline: bool = False, num_blocks: int = 0):
        self.model, self.optimizer = model, optimizer
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.device = model.get_device()
        self.fused = self.device.type == "cuda" and not nccl_baseline
        self.num_blocks = num_blocks
        model.initialize()
        params, grads = model.parameters(), model.gradients()
        self.numels = [p.numel() for p in params]
        offs, total = [], 0
        for n in self.numels:
            offs.append(total)
            total += (n + 3) // 4 * 4
        quantum = 4 * self.world
        self.n = (total + quantum - 1) // quantum * quantum
        self.offsets = offs
        if self.device.type == "cuda":
            # rendezvous state shared by the fused update and sync-BN: symmetric epoch flags + local counters
            self.ctl = SymmetricBuffer(2 * 64, self.group, self.device)
            self.start_flags = self.ctl.peer_array(0)
            self.done_flags = self.ctl.peer_array(64)
            self.local = torch.zeros(8, dtype=torch.int64, device=self.device)  # [0] block ctr, [1] epoch
        self.buckets = None           # DCNN_DP_BUCKETS=K: K fused all-reduce launches overlapped with backward
        self.final_backward = False   # set by the training step before the last backward pass of an optimisation step
        if self.fused:
            self._setup_fused(params)
            import os
            k = int(os.environ.get("DCNN_DP_BUCKETS", "0") or 0)
            if k > 1 and self.world > 1 and not self.nvls:
                self._setup_buckets(k)
        else:
            self.flat_grad = None
            optimizer.attach(params, grads)
        self.sync_objs: List[SyncBN] = []
        if sync_bn and self.world > 1:
            self._setup_sync_bn()
        # replicas must start from identical weights: broadcast rank 0's
        self.broadcast_parameters()

    # ------------------------------------------------------------------------------------------------
    def _setup_fused(self, params):
        n = self.n
        # layout: [params n floats][grads n floats]
        import os
        self.nvls = False
        if os.environ.get("DCNN_DP_NVLS", "0") == "1" and self.world > 1:
            from .._lib import lib
            dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
            ok = torch.tensor([int(lib().dcnn_nvls_supported(C.c_int(dev)))], device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)     # all replicas or none
            self.nvls = bool(int(ok.item()))
            if not self.nvls and self.rank == 0:
                print("DCNN_DP_NVLS=1 but multicast objects are not supported here: using the P2P kernel")
        self.buf = NvlsBuffer(8 * n, self.group, self.device) if self.nvls \
            else SymmetricBuffer(8 * n, self.group, self.device)
        self.flat_param = self.buf.view(torch.float32, 0, n)
        self.flat_grad = self.buf.view(torch.float32, 4 * n, n)
        new_p, new_g = [], []
        for p, off, num in zip(params, self.offsets, self.numels):
            fp = self.flat_param[off:off + num]
            fg = self.flat_grad[off:off + num]
            if p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last) and not p.is_contiguous():
                o, c, r, s = p.shape
                fp = fp.view(o, r, s, c).permute(0, 3, 1, 2)
                fg = fg.view(o, r, s, c).permute(0, 3, 1, 2)
            else:
                fp, fg = fp.view(p.shape), fg.view(p.shape)
            new_p.append(fp)
            new_g.append(fg)
        used = 0
        for l in self.model.layers:
            used += l.rebind_parameters(new_p[used:], new_g[used:])
        assert used == len(new_p)
        shard = self.n // self.world
        self.shard_begin, self.shard_end = self.rank * shard, (self.rank + 1) * shard
        opt = self.optimizer
        self._alloc_state(shard)
        opt.params, opt.grads = self.model.parameters(), self.model.gradients()   # for clear_gradients()/introspection
        # checkpoints: the moments live here, sharded -- the optimizer's state_dict()/load_state_dict() go through us
        opt.state_dict = self.optimizer_state_dict
        opt.load_state_dict = self.load_optimizer_state_dict
        if not self.nvls:
            self.grad_ptrs = self.buf.peer_array(4 * self.n)
            self.param_ptrs = self.buf.peer_array(0)
        # push variant of the fused kernel (default; DCNN_AR_PULL=1 keeps the peer-load kernel): every replica owns a
        # staging buffer that receives the other replicas' slices of its shard
        self.push = (not self.nvls) and self.world > 1 and os.environ.get("DCNN_AR_PULL", "0") != "1"
        if self.push:
            self._setup_push(None)
        torch.cuda.synchronize()
        dist.barrier(group=self.group)

    # ------------------------------------------------------------------------------------------------
    # Sharded optimizer state.  It lives in a symmetric buffer so that ANY rank can assemble the full moments for a
    # checkpoint by reading its peers' shards over NVLink -- no collective, so a signal handler on one rank can do it.
    def _alloc_state(self, floats: int):
        opt = self.optimizer
        self.n_state = (2 if isinstance(opt, Adam) else (1 if getattr(opt, "momentum", 0) > 0 else 0))
        self.state_floats = int(floats)
        self.s1 = self.s2 = None
        if self.n_state:
            self.state_buf = SymmetricBuffer(4 * self.n_state * self.state_floats, self.group, self.device)
            self.s1 = self.state_buf.view(torch.float32, 0, self.state_floats)
            if self.n_state == 2:
                self.s2 = self.state_buf.view(torch.float32, 4 * self.state_floats, self.state_floats)

    def _setup_push(self, buckets):
        """Staging buffer and per-launch tables of the push kernel.  A launch (the whole buffer, or one bucket) covers
        the flat range [a, a+len); replica r owns [sb_r, se_r) of it and receives the other replicas' copies of that
        slice in its staging buffer at (R-1) x (offset of the slice in r's optimizer-state arrays)."""
        R = self.world
        per_rank = [state_layout(self.n, R, r, buckets) for r in range(R)]
        floats = max((R - 1) * sum(e - b for b, e, _ in lay) for lay in per_rank)
        if getattr(self, "stage_buf", None) is None or self.stage_buf.nbytes < 4 * floats:
            self.stage_buf = SymmetricBuffer(4 * max(4, floats), self.group, self.device)
        self.stage_ptrs = self.stage_buf.peer_array(0)
        LLA = C.c_longlong * R
        launches = []
        n_launch = 1 if buckets is None else len(buckets)
        for li in range(n_launch):
            a = 0 if buckets is None else buckets[li]["a"]
            sb = LLA(*[per_rank[r][li][0] - a for r in range(R)])
            se = LLA(*[per_rank[r][li][1] - a for r in range(R)])
            soff = LLA(*[(R - 1) * per_rank[r][li][2] for r in range(R)])
            launches.append((sb, se, soff))
        self.push_tables = launches
        self.push_counters = torch.zeros(2, dtype=torch.int32, device=self.device)

    def _launch_push(self, li, grad_ptrs, param_ptrs, n, s1, s2, args, opt, num_blocks, stream_ptr):
        from .._lib import F, I, LL, P, check, lib
        from ..ops.cuda import _p
        mode, a, b, eps, bc1, bc2, wd = args
        sb, se, soff = self.push_tables[li]
        check(lib().dcnn_fused_allreduce_update_push(
            grad_ptrs, param_ptrs, self.stage_ptrs, sb, se, soff, _p(s1), _p(s2), LL(n), self.start_flags,
            self.done_flags, P(self.push_counters.data_ptr()), I(self.rank), I(self.world),
            P(self.local.data_ptr() + 8), I(mode), F(opt.learning_rate), F(a), F(b), F(eps), F(bc1), F(bc2), F(wd),
            _p(opt.hyper), I(num_blocks), stream_ptr), "dcnn_fused_allreduce_update_push")

    def _state_layout(self, rank: int):
        """[(flat_begin, flat_end, offset in that rank's state arrays)] of the slices rank ``rank`` owns."""
        return state_layout(self.n, self.world, rank, self.buckets)

    def _flat_to_list(self, flat):
        out = []
        for p, off, num in zip(self.model.parameters(), self.offsets, self.numels):
            t = flat[off:off + num]
            if p.dim() == 4 and not p.is_contiguous():
                o, c, r, s_ = p.shape
                t = t.view(o, r, s_, c).permute(0, 3, 1, 2)
            else:
                t = t.view(p.shape)
            out.append(t.clone())
        return out

    def _list_to_flat(self, tensors):
        flat = torch.zeros(self.n, device=self.device)
        for p, off, num, t in zip(self.model.parameters(), self.offsets, self.numels, tensors):
            t = t.to(self.device, torch.float32)
            src = t.permute(0, 2, 3, 1).reshape(-1) if (p.dim() == 4 and not p.is_contiguous()) else t.reshape(-1)
            flat[off:off + num] = src
        return flat

    def _gather_state(self, which: int):
        torch.cuda.synchronize(self.device)
        flat = torch.zeros(self.n, device=self.device)
        for r in range(self.world):
            peer = torch.as_tensor(_RawCuda(self.state_buf.peer_ptrs[r] + 4 * which * self.state_floats,
                                            4 * self.state_floats), device=self.device).view(torch.float32)
            for b, e, off in self._state_layout(r):
                flat[b:e] = peer[off:off + e - b]
        return self._flat_to_list(flat)

    def optimizer_state_dict(self) -> dict:
        """Same layout as the optimizer's own ``state_dict()`` (per-parameter tensors), assembled from every replica's
        shard: a checkpoint written under data parallelism resumes on one GPU and the other way round."""
        opt = self.optimizer
        if isinstance(opt, Adam):
            return {"t": opt.t, "m": self._gather_state(0), "v": self._gather_state(1)}
        return {"velocities": self._gather_state(0) if self.n_state else []}

    def load_optimizer_state_dict(self, state: dict):
        opt = self.optimizer
        if isinstance(opt, Adam):
            opt.t = int(state.get("t", 0))
            pairs = [(self.s1, state.get("m")), (self.s2, state.get("v"))]
        else:
            pairs = [(self.s1, state.get("velocities"))] if self.n_state else []
        for dst, lst in pairs:
            if not lst:
                if dst is not None and lst is not None and len(self.numels):
                    raise ValueError("checkpoint holds no optimizer moments for a stateful optimizer")
                continue
            flat = self._list_to_flat(lst)
            for b, e, off in self._state_layout(self.rank):
                dst[off:off + e - b] = flat[b:e]
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------------------------------------
    # Bucketed, overlapped variant (opt-in, DCNN_DP_BUCKETS=K; written after the last GPU minutes of its session, not yet
    # run on hardware).  The flat gradient buffer is cut at fusion-unit boundaries into K contiguous buckets.  Backward
    # visits units from last to first, so a bucket's gradients are final once its FIRST unit has run its backward; at
    # that moment the same fused kernel is launched for just that range on a side stream (event-ordered after the
    # compute stream), so reduce-scatter + update + all-gather of the late layers run under the backward of the early
    # ones.  Every bucket is sharded over the replicas on its own; the optimizer state of this rank is the
    # concatenation of its bucket shards.
    def _setup_buckets(self, k: int):
        units = self.model._get_units()
        unit_params, i = [], 0          # parameter index range of every unit (model.parameters() is unit order)
        for u in units:
            cnt = sum(len(l.parameters()) for l in u.layers)
            unit_params.append((i, i + cnt))
            i += cnt
        assert i == len(self.offsets), "fusion units do not cover the model parameters"
        self.buckets = plan_buckets(unit_params, self.numels, self.offsets, self.n, k, self.rank, self.world)
        for bk in self.buckets:
            bk["grad_ptrs"] = self.buf.peer_array(4 * self.n + 4 * bk["a"])
            bk["param_ptrs"] = self.buf.peer_array(4 * bk["a"])
        self._alloc_state(self.n // self.world + 4 * len(self.buckets))
        if self.push:
            self._setup_push(self.buckets)
        self.comm_stream = torch.cuda.Stream(device=self.device)
        self._trigger = {bk["first_unit"]: bi for bi, bk in enumerate(self.buckets)}
        self._launched = set()
        self._step_args = None
        self.model.after_unit_backward = self._on_unit_backward

    def _on_unit_backward(self, unit_index: int):
        if self.final_backward and unit_index in self._trigger:
            self._launch_bucket(self._trigger[unit_index])

    def _begin_step(self, opt):
        """Once per optimisation step, before its first bucket launch: step count and bias corrections."""
        if isinstance(opt, Adam):
            opt.t += 1
            bc1, bc2 = opt.bias_corrections()
            args = (3 if opt.decouple_weight_decay else 2, opt.beta1, opt.beta2, opt.epsilon, bc1, bc2, opt.weight_decay)
        elif isinstance(opt, SGD):
            args = (1 if opt.momentum > 0 else 0, opt.momentum, 0.0, 0.0, 1.0, 1.0, 0.0)
        else:
            raise TypeError(f"fused all-reduce+update does not know optimizer {type(opt).__name__}")
        opt.push_hyper()
        self._step_args = args

    def _launch_bucket(self, bi: int):
        from .._lib import F, I, LL, P, check, lib
        from ..ops.cuda import _p
        if bi in self._launched:
            return
        opt = self.optimizer
        if self._step_args is None:
            self._begin_step(opt)
        mode, a, b, eps, bc1, bc2, wd = self._step_args
        bk = self.buckets[bi]
        ready = torch.cuda.Event()
        ready.record()                                   # the bucket's gradients are complete at this point
        self.comm_stream.wait_event(ready)
        s1 = self.s1[bk["state_off"]:] if self.s1 is not None else None
        s2 = self.s2[bk["state_off"]:] if self.s2 is not None else None
        if self.push:
            with torch.cuda.stream(self.comm_stream):
                self._launch_push(bi, bk["grad_ptrs"], bk["param_ptrs"], bk["n"], s1, s2, self._step_args, opt,
                                  self.num_blocks if self.num_blocks > 0 else 32, P(self.comm_stream.cuda_stream))
            self._launched.add(bi)
            return
        with torch.cuda.stream(self.comm_stream):
            check(lib().dcnn_fused_allreduce_update(
                bk["grad_ptrs"], bk["param_ptrs"], _p(s1), _p(s2), LL(bk["n"]), LL(bk["sb"]), LL(bk["se"]),
                self.start_flags, self.done_flags, P(self.local.data_ptr()), I(self.rank), I(self.world),
                P(self.local.data_ptr() + 8), I(mode), F(opt.learning_rate), F(a), F(b), F(eps), F(bc1), F(bc2), F(wd),
                _p(opt.hyper), I(self.num_blocks if self.num_blocks > 0 else 32),
                P(self.comm_stream.cuda_stream)), "dcnn_fused_allreduce_update (bucket)")
        self._launched.add(bi)

    def _finish_buckets(self):
        for bi in range(len(self.buckets) - 1, -1, -1):   # whatever the hooks did not launch (e.g. no final-pass flag)
            self._launch_bucket(bi)
        torch.cuda.current_stream().wait_stream(self.comm_stream)
        self._launched = set()
        self._step_args = None
        self.final_backward = False

    def close(self):
        """Collective: release the symmetric buffers (parameters and gradients move to ordinary tensors first, so the
        model stays usable on its own)."""
        if self.device.type != "cuda":
            return
        if self.fused:
            params = [p.detach().clone(memory_format=torch.preserve_format) for p in self.model.parameters()]
            grads = [g.detach().clone(memory_format=torch.preserve_format) for g in self.model.gradients()]
            used = 0
            for l in self.model.layers:
                used += l.rebind_parameters(params[used:], grads[used:])
            self.flat_param = self.flat_grad = self.s1 = self.s2 = None
            from .. import ops
            ops.weights_changed()
        for l in _all_layers(self.model):
            if isinstance(l, BatchNormLayer):
                l.sync = None
        for name in ("buf", "stage_buf", "state_buf", "bn_buf", "ctl"):
            b = getattr(self, name, None)
            if b is not None:
                b.close()
                setattr(self, name, None)

    def reset_step_bookkeeping(self):
        """Forget a step that was recorded but never ran (an aborted CUDA-graph capture).  The rendezvous epochs and the
        sync-BatchNorm call counters are device-resident, so they did not move; only the bucket bookkeeping is host state."""
        if self.buckets is not None:
            self._launched = set()
            self._step_args = None
        self.final_backward = False

    def _setup_sync_bn(self):
        bns = [l for l in _all_layers(self.model) if isinstance(l, BatchNormLayer)]
        on_gpu = self.device.type == "cuda"
        if on_gpu:
            from .._lib import lib
            l_ = lib()
            l_.dcnn_bn_sync_area_bytes.restype = C.c_longlong
            for l in bns:
                assert SyncBN.area_bytes(self.world, l.num_features) == \
                    int(l_.dcnn_bn_sync_area_bytes(C.c_int(self.world), C.c_int(l.num_features)))
            total = sum(SyncBN.bytes_needed(self.world, l.num_features) for l in bns)
            self.bn_buf = SymmetricBuffer(total, self.group, self.device)     # zero-initialised: epochs start at 0
        off = 0
        for l in bns:
            s = SyncBN(self, l.num_features, off)
            off += SyncBN.bytes_needed(self.world, l.num_features)
            l.sync = s
            self.sync_objs.append(s)
        if on_gpu:
            torch.cuda.synchronize()
            dist.barrier(group=self.group)

    def broadcast_parameters(self, src: int = 0):
        for t in list(self.model.parameters()) + list(self.model.buffers().values()):
            tmp = t.contiguous() if not (t.is_contiguous() or t.dim() != 4) else t
            if t.dim() == 4 and not t.is_contiguous():
                tmp = t.permute(0, 2, 3, 1).contiguous()     # dense O-H-W-I storage order
                dist.broadcast(tmp, src=dist.get_global_rank(self.group, src), group=self.group)
                t.copy_(tmp.permute(0, 3, 1, 2))
            else:
                dist.broadcast(tmp, src=dist.get_global_rank(self.group, src), group=self.group)
        if self.device.type == "cuda":
            torch.cuda.synchronize()

    # ------------------------------------------------------------------------------------------------
    def reduce_and_update(self, optimizer: Optional[Optimizer] = None):
        opt = optimizer or self.optimizer
        if not self.fused:
            return self._baseline(opt)
        from .. import ops
        ops.weights_changed()           # every replica's parameters are rewritten through peer pointers
        if self.buckets is not None:
            return self._finish_buckets()
        from .._lib import F, I, LL, P, check, lib
        from ..ops.cuda import _p, _stream
        if isinstance(opt, Adam):
            opt.t += 1
            bc1, bc2 = opt.bias_corrections()
            mode = 3 if opt.decouple_weight_decay else 2
            a, b, eps, wd = opt.beta1, opt.beta2, opt.epsilon, opt.weight_decay
        elif isinstance(opt, SGD):
            bc1 = bc2 = 1.0
            mode = 1 if opt.momentum > 0 else 0
            a, b, eps, wd = opt.momentum, 0.0, 0.0, 0.0
        else:
            raise TypeError(f"fused all-reduce+update does not know optimizer {type(opt).__name__}")
        opt.push_hyper()
        if self.nvls:
            # reduction inside the NVSwitch (multimem.ld_reduce) and one multicast store per parameter slice
            check(lib().dcnn_fused_allreduce_update_nvls(
                P(self.buf.mc_ptr + 4 * self.n), P(self.buf.mc_ptr), P(self.buf.ptr), P(self.buf.ptr + 4 * self.n),
                _p(self.s1), _p(self.s2), LL(self.n), LL(self.shard_begin), LL(self.shard_end), self.start_flags,
                self.done_flags, P(self.local.data_ptr()), I(self.rank), I(self.world), P(self.local.data_ptr() + 8),
                I(mode), F(opt.learning_rate), F(a), F(b), F(eps), F(bc1), F(bc2), F(wd), _p(opt.hyper),
                I(self.num_blocks), _stream()), "dcnn_fused_allreduce_update_nvls")
            return
        if self.push:
            self._launch_push(0, self.grad_ptrs, self.param_ptrs, self.n, self.s1, self.s2,
                              (mode, a, b, eps, bc1, bc2, wd), opt, self.num_blocks, _stream())
            return
        check(lib().dcnn_fused_allreduce_update(
            self.grad_ptrs, self.param_ptrs, _p(self.s1), _p(self.s2), LL(self.n), LL(self.shard_begin),
            LL(self.shard_end), self.start_flags, self.done_flags, P(self.local.data_ptr()), I(self.rank),
            I(self.world), P(self.local.data_ptr() + 8), I(mode), F(opt.learning_rate), F(a), F(b), F(eps), F(bc1),
            F(bc2), F(wd), _p(opt.hyper), I(self.num_blocks), _stream()), "dcnn_fused_allreduce_update")

    def _baseline(self, opt):
        grads = self.model.gradients()
        if self.device.type == "cuda":
            flat = torch.cat([g.reshape(-1) if g.is_contiguous() else g.permute(0, 2, 3, 1).reshape(-1) for g in grads])
            dist.all_reduce(flat, group=self.group)
            flat.div_(self.world)
            o = 0
            for g in grads:
                n = g.numel()
                if g.is_contiguous():
                    g.copy_(flat[o:o + n].view(g.shape))
                else:
                    oc, c, r, s = g.shape
                    g.copy_(flat[o:o + n].view(oc, r, s, c).permute(0, 3, 1, 2))
                o += n
        else:
            for g in grads:
                dist.all_reduce(g, group=self.group)
                g.div_(self.world)
        opt.update(zero_grads=True)

    def bytes_over_nvlink_per_step(self) -> int:
        """Per-GPU NVLink traffic in ONE direction per step.  P2P kernel: (R-1)/R of the gradients are loaded from the
        peers and (R-1)/R of the parameters arrive from them: 2 * 4n * (R-1)/R bytes inbound (and the same outbound).
        NVLS kernel: 1/R of the gradients arrives already reduced by the switch and (R-1)/R of the parameters arrive by
        multicast: 4n bytes inbound."""
        if getattr(self, "nvls", False):
            return int(self.n * 4)
        return int(2 * self.n * 4 * (self.world - 1) / self.world)
