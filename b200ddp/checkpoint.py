"""Checkpoint I/O, file-compatible with the reference in both directions.

Format (reference ``train.py:178-209``; ground truth in SURVEY §2.6): a
``torch.save`` zip of ``{"epoch": int, "model_state_dict", "optimizer_state_dict",
"loss": float}``, model keys un-prefixed (``layers.0.weight`` …), every tensor
fp32 with its own storage, optimizer state in ``torch.optim`` layout.

Differences from the reference, none visible in the file:

* tensors are exported from the flat arenas as per-tensor clones (the runtime
  layout is one shared arena; the file must not carry it);
* the write is atomic (temp file + ``os.replace``) so a rank killed mid-save
  never leaves a truncated ``latest_model.pt``;
* ``load_checkpoint`` maps to CPU first and copies into the arenas, so resuming
  does not allocate a second copy of the model on the device.
"""
from __future__ import annotations

import os
from typing import Any, Dict

import torch

from .utils.logging import get_logger

logger = get_logger("b200ddp")


def _model_state(model) -> Dict[str, torch.Tensor]:
    if hasattr(model, "model_state_dict"):
        sd = model.model_state_dict()
    else:
        inner = model.module if hasattr(model, "module") else model
        sd = {k: v.detach().clone() for k, v in inner.state_dict().items()}
    return {k: v.cpu() if v.device.type != "cpu" else v for k, v in sd.items()}


def _to_cpu(obj: Any) -> Any:
    if torch.is_tensor(obj):
        return obj.detach().cpu()
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def save_checkpoint(model, optimizer, epoch: int, loss: float, path: str) -> None:
    checkpoint = {
        "epoch": int(epoch),
        "model_state_dict": _model_state(model),
        "optimizer_state_dict": _to_cpu(optimizer.state_dict()),
        "loss": float(loss),
    }
    directory = os.path.dirname(os.path.abspath(path))
    os.makedirs(directory, exist_ok=True)
    tmp = f"{path}.tmp{os.getpid()}"
    try:
        torch.save(checkpoint, tmp)
        os.replace(tmp, path)
    finally:
        if os.path.exists(tmp):
            os.unlink(tmp)
    logger.info(f"Checkpoint saved to {path}")


def load_checkpoint(model, optimizer, path: str, device: torch.device) -> int:
    """Loads ``path`` into ``model``/``optimizer``; returns the stored epoch.

    The stored value is returned unchanged, as the reference does
    (``train.py:209``) — a run resumed from the checkpoint of epoch *e* therefore
    starts *at* epoch *e* unless the caller opts into ``--resume-next-epoch``.
    """
    checkpoint = torch.load(path, map_location="cpu", weights_only=False)
    sd = checkpoint["model_state_dict"]
    if hasattr(model, "load_model_state_dict"):
        model.load_model_state_dict(sd)
    elif hasattr(model, "module"):
        model.module.load_state_dict(sd)
    else:
        model.load_state_dict(sd)
    optimizer.load_state_dict(checkpoint["optimizer_state_dict"])
    logger.info(f"Checkpoint loaded from {path}, epoch {checkpoint['epoch']}")
    return int(checkpoint["epoch"])
