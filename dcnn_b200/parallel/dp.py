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


class SyncBN:
    """Cross-replica BatchNorm statistics for one BN layer (forward moments and backward sums)."""

    def __init__(self, dp: "DataParallel", channels: int, offset_bytes: int):
        self.dp, self.C = dp, channels
        self.off = offset_bytes          # [fwd parity0, fwd parity1, bwd parity0, bwd parity1] x 3*C doubles
        self.calls_f = self.calls_b = 0

    @staticmethod
    def bytes_needed(channels: int) -> int:
        return 4 * 3 * channels * 8

    def _slot(self, which: int, parity: int) -> int:
        return self.off + (which * 2 + parity) * 3 * self.C * 8

    # ---- CUDA: export local moments, then rendezvous + merge inside the finalize kernel
    def finalize_forward(self, stats, eps, momentum, rm, rv, gamma, beta, mean, invstd, scale, shift):
        from .._lib import F, I, P, check, lib
        from ..ops.cuda import _p, _stream
        dp = self.dp
        off = self._slot(0, self.calls_f & 1)
        self.calls_f += 1
        check(lib().dcnn_bn_finalize(_p(stats.partials), I(stats.tiles), I(self.C), F(eps), F(momentum), P(0), P(0),
                                     _p(gamma), _p(beta), _p(mean), _p(invstd), _p(scale), _p(shift),
                                     P(dp.bn_buf.ptr + off), _stream()), "bn_finalize(export)")
        check(lib().dcnn_bn_finalize_sync(dp.bn_buf.peer_array(off), dp.bn_flags, I(dp.rank), P(dp.bn_epoch_ptr),
                                          I(dp.world), I(self.C), F(eps), F(momentum), _p(rm), _p(rv), _p(gamma),
                                          _p(beta), _p(mean), _p(invstd), _p(scale), _p(shift), _stream()),
              "bn_finalize_sync")

    def finalize_backward(self, partials, tiles, Cc, M, mean, invstd, gamma, dgamma, dbeta, k1, k2, k3):
        from .._lib import D, I, P, check, lib
        from ..ops.cuda import _p, _stream
        dp = self.dp
        off = self._slot(1, self.calls_b & 1)
        self.calls_b += 1
        check(lib().dcnn_bn_bwd_finalize(_p(partials), I(tiles), I(Cc), D(float(M)), _p(mean), _p(invstd), _p(gamma),
                                         _p(dgamma), _p(dbeta), _p(k1), _p(k2), _p(k3), P(dp.bn_buf.ptr + off),
                                         _stream()), "bn_bwd_finalize(export)")
        check(lib().dcnn_bn_bwd_finalize_sync(dp.bn_buf.peer_array(off), dp.bn_flags, I(dp.rank),
                                              P(dp.bn_epoch_ptr), I(dp.world), I(Cc), _p(mean), _p(invstd),
                                              _p(gamma), _p(k1), _p(k2), _p(k3), _stream()), "bn_bwd_finalize_sync")

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
            self.ctl = SymmetricBuffer(3 * 64, self.group, self.device)
            self.start_flags = self.ctl.peer_array(0)
            self.done_flags = self.ctl.peer_array(64)
            self.bn_flags = self.ctl.peer_array(128)
            self.local = torch.zeros(8, dtype=torch.int64, device=self.device)  # [0] block ctr, [1] epoch, [2] bn epoch
            self.bn_epoch_ptr = self.local.data_ptr() + 16
        if self.fused:
            self._setup_fused(params)
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
        self.buf = SymmetricBuffer(8 * n, self.group, self.device)
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
        self.s1 = torch.zeros(shard, device=self.device) if (isinstance(opt, Adam) or getattr(opt, "momentum", 0) > 0) \
            else None
        self.s2 = torch.zeros(shard, device=self.device) if isinstance(opt, Adam) else None
        opt.params, opt.grads = self.model.parameters(), self.model.gradients()   # for clear_gradients()/introspection
        self.grad_ptrs = self.buf.peer_array(4 * self.n)
        self.param_ptrs = self.buf.peer_array(0)
        torch.cuda.synchronize()
        dist.barrier(group=self.group)

    def _setup_sync_bn(self):
        bns = [l for l in _all_layers(self.model) if isinstance(l, BatchNormLayer)]
        on_gpu = self.device.type == "cuda"
        if on_gpu:
            total = sum(SyncBN.bytes_needed(l.num_features) for l in bns)
            self.bn_buf = SymmetricBuffer(total, self.group, self.device)
        off = 0
        for l in bns:
            s = SyncBN(self, l.num_features, off)
            off += SyncBN.bytes_needed(l.num_features)
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
        """Per-GPU one-direction NVLink traffic of the fused kernel: (R-1)/R of the gradients in, the same amount of
        parameters out."""
        return int(self.n * 4 * (self.world - 1) / self.world)
