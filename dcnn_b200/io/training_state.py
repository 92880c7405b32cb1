"""True resume (the reference cannot: it persists neither BN running statistics, optimizer moments/step, learning
rate, epoch nor RNG state -- SURVEY §5.4).  ``save_training_state`` writes the reference-compatible ``<path>.json`` /
``<path>.bin`` model snapshot plus ``<path>.state.{json,bin}`` (BN buffers) and ``<path>.train.pt`` (optimizer state,
step counters, schedule position, epoch, RNG states); the reference's loader ignores the extra files."""
from __future__ import annotations

from typing import Optional

import torch


def _plain(obj):
    """numpy arrays / scalars -> tensors / Python numbers, recursively, so that the blob loads with weights_only=True."""
    import numpy as np
    if isinstance(obj, np.ndarray):
        return {"__ndarray__": torch.from_numpy(np.ascontiguousarray(obj).copy())}
    if isinstance(obj, np.generic):
        return obj.item()
    if isinstance(obj, dict):
        return {k: _plain(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [_plain(v) for v in obj]
    return obj


def _unplain(obj):
    if isinstance(obj, dict):
        if set(obj) == {"__ndarray__"}:
            return obj["__ndarray__"].numpy()
        return {k: _unplain(v) for k, v in obj.items()}
    if isinstance(obj, list):
        return [_unplain(v) for v in obj]
    return obj


def save_training_state(path: str, model, optimizer, epoch: int = 0, scheduler=None, extra: Optional[dict] = None):
    model.save_to_file(path, save_state=True)
    opt_state = {k: ([t.detach().cpu() for t in v] if isinstance(v, list) else v)
                 for k, v in optimizer.state_dict().items()}
    blob = {"epoch": int(epoch), "optimizer_config": optimizer.get_config(), "optimizer_state": opt_state,
            "learning_rate": optimizer.get_learning_rate(),
            "scheduler": None if scheduler is None else {"config": scheduler.get_config(),
                                                          "current_step": scheduler.get_current_step()},
            "torch_rng": torch.get_rng_state(),
            "cuda_rng": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None,
            "extra": _plain(extra or {})}
    torch.save(blob, path + ".train.pt")


def load_training_state(path: str, model, optimizer, scheduler=None) -> dict:
    """Restores weights, BN statistics, optimizer state (the optimizer must already be attached), LR and RNG.
    Returns the saved dict (``epoch``, ``extra`` ...)."""
    model.load_weights_file(path + ".bin")
    blob = torch.load(path + ".train.pt", weights_only=True)   # tensors, numbers, dicts, lists only: no code runs
    blob["extra"] = _unplain(blob.get("extra", {}))
    optimizer.load_state_dict(blob["optimizer_state"])
    optimizer.set_learning_rate(blob["learning_rate"])
    if scheduler is not None and blob.get("scheduler"):
        scheduler.current_step = int(blob["scheduler"]["current_step"])
    torch.set_rng_state(blob["torch_rng"])
    if blob.get("cuda_rng") is not None and torch.cuda.is_available():
        try:
            torch.cuda.set_rng_state_all(blob["cuda_rng"])
        except RuntimeError:
            pass
    return blob
