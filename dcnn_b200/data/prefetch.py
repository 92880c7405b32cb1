"""Streaming, prefetching front end for any data loader.

The reference's loaders (include/data_loading/data_loader.hpp:120-187) materialise every batch of the epoch in
``prepare_batches`` -- augmented once, all resident -- and the training loop copies each batch to the device when it
needs it.  For a Tiny-ImageNet-sized set that is ~5 GB of pinned host memory and an augmentation that never changes
between epochs.  ``PrefetchLoader`` keeps the loader interface the training loops use (``shuffle / reset /
get_next_batch / num_batches / size``) but builds batches on a worker thread, ``depth`` batches ahead of the consumer:
sample gather and the per-sample augmentation loops run in the native host runtime (they release the GIL), each batch
lands in a reused pinned buffer, and -- when a CUDA device is given -- its host-to-device copy is issued on a side
stream so that it overlaps the previous step's kernels.  Every epoch draws fresh augmentation parameters.

The batch order and contents are those of ``loader.get_batch(batch_size)`` called in sequence (same shuffle, same
sharding, same augmentation draws), so a run with and without the wrapper sees identical data."""
from __future__ import annotations

import queue
import threading
from typing import Optional

import torch


class PrefetchLoader:
    def __init__(self, loader, batch_size: int, depth: int = 3, device=None, drop_last: bool = False):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.loader = loader
        self.batch_size = int(batch_size)
        self.depth = int(depth)
        self.drop_last = bool(drop_last)
        self.device = torch.device(device) if device is not None else None
        self._cuda = self.device is not None and self.device.type == "cuda"
        self._stream = torch.cuda.Stream(self.device) if self._cuda else None
        self._q: Optional[queue.Queue] = None
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self._error: Optional[BaseException] = None
        self._pinned = []          # ring of reusable pinned staging buffers, depth + 2 deep
        self._slot = 0
        self.batches_produced = 0

    # -- loader interface ----------------------------------------------------------------------------------------
    def size(self) -> int:
        return self.loader.size()

    def num_batches(self) -> int:
        n = self.loader.shard_size() if hasattr(self.loader, "shard_size") else self.loader.size()
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def get_batch_size(self) -> int:
        return self.batch_size

    def are_batches_prepared(self) -> bool:
        return True

    def prepare_batches(self, batch_size: int):
        """Kept for interface compatibility: only records the batch size (nothing is materialised)."""
        self._halt()
        self.batch_size = int(batch_size)

    def set_shard(self, rank: int, world: int):
        self._halt()
        self.loader.set_shard(rank, world)

    def set_seed(self, seed: int):
        self.loader.set_seed(seed)

    def set_augmentation(self, aug):
        self._halt()
        self.loader.set_augmentation(aug)

    def get_rng_state(self):
        self._halt()
        return self.loader.get_rng_state()

    def set_rng_state(self, state):
        self._halt()
        self.loader.set_rng_state(state)

    def __getattr__(self, name):          # image shape, class names, ... come from the wrapped loader
        return getattr(self.loader, name)

    def shuffle(self):
        self._halt()
        self.loader.shuffle()

    def reset(self):
        self._halt()
        self.loader.reset()

    def get_next_batch(self):
        """-> (data, labels) on ``device`` (or pinned host tensors), or None at the end of the epoch."""
        if self._thread is None:
            self._start()
        item = self._q.get()
        if item is None:
            self._thread.join()
            self._thread = None
            if self._error is not None:
                err, self._error = self._error, None
                raise err
            return None
        data, labels, event = item
        if event is not None:
            torch.cuda.current_stream(self.device).wait_event(event)
            # the tensors were allocated on the side stream: tell the caching allocator who uses them now
            data.record_stream(torch.cuda.current_stream(self.device))
            labels.record_stream(torch.cuda.current_stream(self.device))
        return data, labels

    def __iter__(self):
        self.reset()
        while True:
            b = self.get_next_batch()
            if b is None:
                return
            yield b

    def close(self):
        self._halt()

    # -- producer ------------------------------------------------------------------------------------------------
    def _staging(self, like: torch.Tensor, which: int) -> torch.Tensor:
        """A pinned buffer of ``like``'s shape from the ring (the consumer is at most depth + 1 batches behind)."""
        ring = self.depth + 2
        while len(self._pinned) < ring:
            self._pinned.append([None, None])
        cur = self._pinned[self._slot % ring][which]
        if cur is None or cur.shape != like.shape or cur.dtype != like.dtype:
            cur = torch.empty(like.shape, dtype=like.dtype)
            if torch.cuda.is_available():
                try:
                    cur = cur.pin_memory()
                except RuntimeError:
                    pass
            self._pinned[self._slot % ring][which] = cur
        return cur

    def _produce(self):
        try:
            while not self._stop.is_set():
                b = self.loader.get_batch(self.batch_size)
                if b is None or (self.drop_last and b[0].shape[0] < self.batch_size):
                    break
                d, l = b
                if self._cuda:
                    hd, hl = self._staging(d, 0).copy_(d), self._staging(l, 1).copy_(l)
                    self._slot += 1
                    with torch.cuda.stream(self._stream):
                        dd = hd.to(self.device, non_blocking=True)
                        dl = hl.to(self.device, non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(self._stream)
                    item = (dd, dl, ev)
                else:
                    item = (d, l, None)
                self.batches_produced += 1
                while not self._stop.is_set():
                    try:
                        self._q.put(item, timeout=0.05)
                        break
                    except queue.Full:
                        continue
        except BaseException as e:      # surfaces in the consumer at the end of the stream
            self._error = e
        finally:
            while True:
                try:
                    self._q.put(None, timeout=0.05)
                    break
                except queue.Full:
                    if self._stop.is_set():
                        try:
                            self._q.get_nowait()
                        except queue.Empty:
                            pass

    def _start(self):
        self._stop.clear()
        self._q = queue.Queue(maxsize=self.depth)
        self._thread = threading.Thread(target=self._produce, name="dcnn-prefetch", daemon=True)
        self._thread.start()

    def _halt(self):
        """Stop the producer (if any) and drop what it had queued."""
        if self._thread is None:
            return
        self._stop.set()
        while self._thread.is_alive():
            try:
                self._q.get(timeout=0.05)
            except queue.Empty:
                pass
        self._thread.join()
        self._thread = None
        self._q = None

    def __del__(self):
        try:
            self._halt()
        except Exception:
            pass
