"""Running-average meter and top-1 accuracy (reference utils.py:5-27)."""
from __future__ import annotations

import torch


class AverageMeter:
    """``update(val, n)`` keeps val / sum / count / avg exactly like utils.py:5-20.

    Values may be Python numbers or tensors.  Tensors are accumulated ON DEVICE (no
    implicit ``.item()``), so feeding the per-step loss costs no host sync; call
    ``.item()``/format only when a number is actually printed."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.val = 0
        self.avg = 0
        self.sum = 0
        self.count = 0

    def update(self, val, n=1):
        if torch.is_tensor(val):
            val = val.detach()
        self.val = val
        self.sum = self.sum + val * n
        self.count += n
        self.avg = self.sum / self.count

    def values(self):
        """(val, avg) as Python floats -- the one place a device sync happens."""
        f = lambda v: float(v.item()) if torch.is_tensor(v) else float(v)
        return f(self.val), f(self.avg)


def Accuracy(out: torch.Tensor, label: torch.Tensor) -> torch.Tensor:
    """Per-sample top-1 hit as an int64 tensor of shape [N] (utils.py:25-27: ``1 * (pred == label)``)."""
    from .. import ops
    return ops.top1_hits(out, label)
