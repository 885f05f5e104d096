"""On-device synthetic image source (replaces the disk pipeline that cannot run offline).

Every batch is generated AND normalised by one kernel directly in NHWC bf16 on the
GPU (no H2D copy, no separate normalise pass).  Pixel values are a
pure function of (seed, rank-unique step counter, linear index), so runs are
reproducible and ranks see different data, like a sharded dataset.  On CPU the torch
oracle produces bit-identical values.

``SyntheticHostLoader`` is the end-to-end variant: uint8 batches live in PINNED host
memory and every step pays a real H2D copy + device normalise, which is what
``bench.py`` times for its ``e2e`` number.
"""
from __future__ import annotations

from typing import Iterator, Tuple

import torch

from .. import ops
from ..parallel.dist import get_rank, get_world_size


class SyntheticDeviceLoader:
    def __init__(self, batch_size: int, image_size: int, num_classes: int, steps_per_epoch: int,
                 device="cpu", dtype=None, seed: int = 0, channels: int = 3, rank=None, world=None):
        self.batch_size, self.image_size, self.num_classes = batch_size, image_size, num_classes
        self.steps_per_epoch, self.device, self.seed, self.channels = steps_per_epoch, torch.device(device), seed, channels
        self.dtype = dtype or (torch.bfloat16 if self.device.type == "cuda" else torch.float32)
        self.rank = get_rank() if rank is None else rank
        self.world = get_world_size() if world is None else world
        self.epoch = 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __len__(self):
        return self.steps_per_epoch

    def batch(self, step: int) -> Tuple[torch.Tensor, torch.Tensor]:
        gstep = (self.epoch * self.steps_per_epoch + step) * self.world + self.rank
        return ops.synthetic_batch(self.batch_size, self.channels, self.image_size, self.image_size,
                                   self.num_classes, self.seed, gstep, self.device, self.dtype)

    def __iter__(self) -> Iterator[dict]:
        for s in range(self.steps_per_epoch):
            img, lab = self.batch(s)
            yield {"image": img, "label": lab, "image_id": [f"syn_{self.epoch}_{s}_{i}" for i in range(self.batch_size)]}


class SyntheticHostLoader:
    """uint8 NHWC batches in pinned host memory -> async H2D -> device normalise.

    ``n_buffers`` distinct host batches are cycled so the copy source changes every step."""

    def __init__(self, batch_size: int, image_size: int, num_classes: int, steps_per_epoch: int,
                 device="cuda", seed: int = 0, n_buffers: int = 4, dtype=torch.bfloat16):
        self.batch_size, self.image_size, self.num_classes = batch_size, image_size, num_classes
        self.steps_per_epoch, self.device, self.dtype = steps_per_epoch, torch.device(device), dtype
        g = torch.Generator().manual_seed(seed + 7919 * get_rank())
        pin = self.device.type == "cuda"
        self.images = [torch.randint(0, 256, (batch_size, image_size, image_size, 3), dtype=torch.uint8,
                                     generator=g) for _ in range(n_buffers)]
        self.labels = [torch.randint(0, num_classes, (batch_size,), dtype=torch.int64, generator=g)
                       for _ in range(n_buffers)]
        if pin:
            self.images = [t.pin_memory() for t in self.images]
            self.labels = [t.pin_memory() for t in self.labels]

    @property
    def h2d_bytes_per_step(self) -> int:
        return self.images[0].numel() * self.images[0].element_size() + self.labels[0].numel() * 8

    def __len__(self):
        return self.steps_per_epoch

    def host_batch(self, step: int) -> dict:
        """The pinned uint8 batch of this step, still on the host (what a DataLoader with pin_memory yields)."""
        i = step % len(self.images)
        return {"image": self.images[i], "label": self.labels[i], "image_id": None}

    def batch(self, step: int):
        i = step % len(self.images)
        img = self.images[i].to(self.device, non_blocking=True)
        lab = self.labels[i].to(self.device, non_blocking=True)
        return ops.normalize_u8(img, self.dtype), lab

    def host_iter(self, steps=None):
        for s in range(self.steps_per_epoch if steps is None else steps):
            yield self.host_batch(s)

    def __iter__(self):
        for s in range(self.steps_per_epoch):
            img, lab = self.batch(s)
            yield {"image": img, "label": lab, "image_id": None}
