"""Host -> device prefetch on a copy stream.

The reference moves each batch with ``.cuda(non_blocking=True)`` on the compute stream right before the forward pass
(train.py:45-46), so the PCIe copy of step i sits between the kernels of step i-1 and step i.  ``DevicePrefetcher``
issues the copy (and the uint8 -> bf16 normalisation of raw loaders) of batch i+1 on a second stream while step i
computes; the consumer only waits on an event.  Works for any iterable of ``{"image", "label", ...}`` dicts."""
from __future__ import annotations

from typing import Iterable, Iterator, Optional

import torch

from .. import ops


class DevicePrefetcher:
    def __init__(self, loader: Iterable, device, dtype: Optional[torch.dtype] = None):
        self.loader, self.device = loader, torch.device(device)
        self.dtype = dtype or (torch.bfloat16 if self.device.type == "cuda" else torch.float32)
        self.stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None

    def __len__(self):
        return len(self.loader)

    def __getattr__(self, name):            # sampler / set_epoch / dataset of the wrapped loader stay reachable
        return getattr(self.loader, name)

    def _stage(self, batch):
        """Start moving one host batch; returns (batch on device, event or None)."""
        if self.stream is None:
            return batch, None
        with torch.cuda.stream(self.stream):
            out = dict(batch)
            img, lab = batch["image"], batch["label"]
            if not torch.is_tensor(lab):
                lab = torch.as_tensor(lab)
            img = img.to(self.device, non_blocking=True)
            lab = lab.to(self.device, non_blocking=True)
            if img.dtype == torch.uint8:
                img = ops.normalize_u8(img, self.dtype)
            out["image"], out["label"] = img, lab
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return out, ev

    def _hand_over(self, staged):
        batch, ev = staged
        if ev is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            for k in ("image", "label"):
                batch[k].record_stream(cur)      # the allocator must not recycle these while the step still reads them
        return batch

    def __iter__(self) -> Iterator[dict]:
        it = iter(self.loader)
        try:
            staged = self._stage(next(it))
        except StopIteration:
            return
        for nxt in it:
            ready = self._hand_over(staged)
            staged = self._stage(nxt)            # overlaps the step that consumes `ready`
            yield ready
        yield self._hand_over(staged)
