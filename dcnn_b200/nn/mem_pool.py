"""Size-keyed pool of free tensors and a scoped temporary buffer (ref include/nn/mem_pool.hpp:11-101).

The reference's layers draw their scratch tensors from a ``MemPool`` (a multimap of free blocks keyed by capacity, best
fit on ``get``, guarded by a mutex) through the RAII ``TempBuffer``.  On CUDA the PyTorch caching allocator and the
CUDA-graph memory pools already do that job for the hot path, so the layers here do not route through this class; it
exists for user code written against the reference's API and for host-side scratch that should be recycled across
steps (pinned staging buffers are expensive to allocate)."""
from __future__ import annotations

import bisect
import threading
from typing import List, Sequence, Tuple

import torch


class MemPool:
    def __init__(self, dtype: torch.dtype = torch.float32, pin_memory: bool = False):
        self.dtype = dtype
        self.pin_memory = pin_memory
        self._lock = threading.Lock()
        self._free: List[Tuple[int, int, torch.Tensor]] = []   # (capacity in elements, serial, flat tensor), sorted
        self._serial = 0
        self.hits = 0
        self.misses = 0

    @staticmethod
    def _numel(shape: Sequence[int]) -> int:
        n = 1
        for d in shape:
            n *= int(d)
        return n

    def get(self, shape: Sequence[int], device="cpu") -> torch.Tensor:
        """A tensor of ``shape`` on ``device``: the smallest free block that is large enough, else a new allocation.
        Contents are unspecified."""
        device = torch.device(device)
        need = self._numel(shape)
        with self._lock:
            i = bisect.bisect_left(self._free, (need, -1))
            while i < len(self._free):
                cap, _, block = self._free[i]
                if block.device == device:
                    del self._free[i]
                    self.hits += 1
                    out = block[:need].view(*shape)
                    out._mempool_block = block          # keeps the whole block reachable for release()
                    return out
                i += 1
            self.misses += 1
        block = torch.empty(max(need, 1), dtype=self.dtype, device=device,
                            pin_memory=self.pin_memory and device.type == "cpu" and torch.cuda.is_available())
        out = block[:need].view(*shape)
        out._mempool_block = block
        return out

    def release(self, tensor: torch.Tensor):
        """Give a tensor obtained from ``get`` (or any contiguous tensor of this pool's dtype) back to the pool."""
        block = getattr(tensor, "_mempool_block", None)
        if block is None:
            if tensor.dtype != self.dtype or not tensor.is_contiguous():
                raise ValueError("MemPool.release: tensor does not belong to this pool")
            block = tensor.reshape(-1)
        with self._lock:
            self._serial += 1
            bisect.insort(self._free, (block.numel(), self._serial, block))

    def clear(self):
        with self._lock:
            self._free.clear()

    def free_blocks(self) -> int:
        with self._lock:
            return len(self._free)

    def free_bytes(self) -> int:
        with self._lock:
            return sum(b.numel() * b.element_size() for _, _, b in self._free)


class TempBuffer:
    """Scoped buffer from a pool (ref TempBuffer): ``with TempBuffer(pool, shape, device) as t: ...`` returns the block
    on exit; ``get()`` / ``release()`` for manual scoping."""

    def __init__(self, pool: MemPool, shape: Sequence[int], device="cpu"):
        self._pool = pool
        self._tensor = pool.get(shape, device)

    def get(self) -> torch.Tensor:
        if self._tensor is None:
            raise RuntimeError("TempBuffer already released")
        return self._tensor

    def release(self):
        if self._tensor is not None:
            self._pool.release(self._tensor)
            self._tensor = None

    def __enter__(self) -> torch.Tensor:
        return self.get()

    def __exit__(self, *exc):
        self.release()
        return False

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass
