"""Checkpoint save / resume with the reference's on-disk layout.

Layout (train.py:176-180): ``torch.save({'epoch', 'best_score', 'state_dict'})`` where the
state dict carries DDP's ``module.`` prefix (``module.encoder.conv1.weight`` ...);
``best_model`` on improvement and ``latest_model`` every 5th epoch under
``<ckpt_dir>/<name>/`` (default ``dtmodel/cp``).  Resume is the reference's
key-intersection merge (train.py:136-150).  Fixes: directories are created (D9) and
``start_epoch`` is returned so the driver can honour it (D10).  Optional extra keys
(optimizer / scheduler / sampler epoch) are additive and ignored by the reference.
"""
from __future__ import annotations

import os
from typing import Any, Dict, Optional, Tuple

import torch

BEST, LATEST = "best_model", "latest_model"


def checkpoint_path(ckpt_dir: str, name: str, which: str = BEST) -> str:
    return os.path.join(ckpt_dir, name, which)


def prefixed_state_dict(model: torch.nn.Module, prefix: str = "module.") -> Dict[str, torch.Tensor]:
    """State dict with the ``module.`` prefix the reference's DDP wrapper produces.
    Tensors are returned in standard contiguous layout on the CPU."""
    inner = getattr(model, "module", model)
    return {prefix + k: v.detach().to("cpu").contiguous() for k, v in inner.state_dict().items()}


def save_checkpoint(model, epoch: int, best_score: float, ckpt_dir: str, name: str, which: str = BEST,
                    extra: Optional[Dict[str, Any]] = None) -> str:
    path = checkpoint_path(ckpt_dir, name, which)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    blob = {"epoch": epoch, "best_score": float(best_score), "state_dict": prefixed_state_dict(model)}
    if extra:
        blob.update(extra)
    tmp = path + ".tmp"
    torch.save(blob, tmp)
    os.replace(tmp, path)            # atomic: a crash never leaves a truncated best_model
    return path


def load_checkpoint(model, ckpt_dir: str, name: str, which: str = BEST) -> Tuple[int, float, Dict[str, Any]]:
    """Returns (start_epoch, best_score, blob); (0, 0, {}) if no checkpoint exists.
    Every key present in both dicts (with or without the ``module.`` prefix) and of equal
    shape is copied; the rest keep their current values (silent partial load, as upstream)."""
    path = checkpoint_path(ckpt_dir, name, which)
    if not os.path.exists(path):
        return 0, 0.0, {}
    blob = torch.load(path, map_location="cpu", weights_only=False)
    loaded = blob["state_dict"]
    inner = getattr(model, "module", model)
    own = inner.state_dict()
    merged = {}
    for k, v in own.items():
        src = loaded.get("module." + k, loaded.get(k))
        merged[k] = src.to(v.dtype) if (src is not None and src.shape == v.shape) else v
    inner.load_state_dict(merged)
    return int(blob.get("epoch", 0)), float(blob.get("best_score", 0.0)), blob
