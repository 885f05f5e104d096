"""Folder-of-classes image dataset, augmentation and loaders (reference dp/loader.py).

Kept from the reference: ``<datadir>/{train,valid}/<class>/*`` layout (dp/loader.py:20-21),
first three channels + nearest-neighbour resize (:44-45), train-only augmentation of
rot90 / flips / one-of saturation-brightness-contrast with factor 0.9+U*0.2 (:63-83),
ImageNet normalisation (:86-91) and the sample dict ``{'image','label','image_id'}`` (:59-61).

Fixed: the label mapping is built from the sorted class directories (D5),
the file order is a *seeded* shuffle identical on every rank so sharding is a true
partition (D12), and the three colour functions that upstream imports from a missing
module are restated here (D3).

B200-first addition: ``raw_u8=True`` skips CPU normalisation and yields uint8 HWC so the
batch crosses PCIe at 1 byte/px and is normalised + cast to NHWC bf16 by a device kernel
(:func:`picd_b200.ops.normalize_u8`), replacing the 3-iteration Python loop and K21.
"""
from __future__ import annotations

import os
import random
from pathlib import Path
from typing import Dict, List, Optional

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, get_worker_info

from ..ops.reference import IMAGENET_MEAN, IMAGENET_STD
from ..parallel.sampler import ShardedSampler

try:  # cv2 is optional; PIL covers decode + nearest resize when it is absent
    import cv2
except Exception:  # pragma: no cover
    cv2 = None


# ---- colour ops (restated; upstream's come from the absent bs.dp.augumentation_utils) ----
def _clip_u8(x):
    return np.clip(x, 0, 255).astype(np.uint8)


def saturation(image: np.ndarray, factor: float) -> np.ndarray:
    gray = image.astype(np.float32) @ np.array([0.299, 0.587, 0.114], dtype=np.float32)
    return _clip_u8(gray[..., None] + factor * (image.astype(np.float32) - gray[..., None]))


def brightness(image: np.ndarray, factor: float) -> np.ndarray:
    return _clip_u8(image.astype(np.float32) * factor)


def contrast(image: np.ndarray, factor: float) -> np.ndarray:
    mean = image.astype(np.float32).mean()
    return _clip_u8(mean + factor * (image.astype(np.float32) - mean))


def _imread_rgb(path: str) -> np.ndarray:
    if cv2 is not None:
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        if img is None:
            raise OSError(f"cannot decode {path}")
        if img.ndim == 2:
            img = np.stack([img] * 3, -1)
        img = img[..., :3][..., ::-1]          # BGR(A) -> RGB
        return np.ascontiguousarray(img)
    from PIL import Image
    return np.asarray(Image.open(path).convert("RGB"))


def _resize_nearest(img: np.ndarray, size: int) -> np.ndarray:
    if cv2 is not None:
        return cv2.resize(img, (size, size), interpolation=cv2.INTER_NEAREST)
    from PIL import Image
    return np.asarray(Image.fromarray(img).resize((size, size), Image.NEAREST))


class ImageDataset(Dataset):
    def __init__(self, data_dir: str, fold: str, resize_size: int, seed: int = 0, raw_u8: bool = False,
                 class_names: Optional[List[str]] = None):
        super().__init__()
        self.data_dir, self.fold, self.resize_size, self.raw_u8 = data_dir, fold, resize_size, raw_u8
        self.image_dir = Path(data_dir) / fold
        files = sorted(str(p) for p in self.image_dir.glob("*/*") if p.is_file())
        random.Random(seed).shuffle(files)       # same order on every rank (fix D12)
        self.image_files = files
        names = class_names if class_names is not None else sorted(
            d.name for d in self.image_dir.iterdir() if d.is_dir()) if self.image_dir.is_dir() else []
        self.mapping: Dict[str, int] = {n: i for i, n in enumerate(names)}   # fix D5
        self.seed = seed
        self._rng: Optional[random.Random] = None     # created lazily, once per process: see _augment_rng
        self._rng_owner = None

    def _augment_rng(self) -> random.Random:
        """Augmentation RNG of THIS process.  DataLoader workers are forks of the parent, so an RNG created in __init__
        would be copied into every worker and all of them would draw the same rot90 / flip / colour decisions.  torch
        gives each worker a distinct seed (base_seed + worker id, re-drawn per epoch from the global generator that
        ``--seed`` seeds), which is what the reference's use of the global ``random`` module relies on; the main process
        (num_workers=0) derives its stream from ``seed``."""
        info = get_worker_info()
        key = (os.getpid(), None if info is None else info.seed)
        if self._rng is None or self._rng_owner != key:
            self._rng = random.Random(self.seed * 1000003 + 17 if info is None else info.seed)
            self._rng_owner = key
        return self._rng

    def __len__(self):
        return len(self.image_files)

    @property
    def num_classes(self) -> int:
        return len(self.mapping)

    def __getitem__(self, idx):
        path = self.image_files[idx]
        image_id = os.path.splitext(os.path.basename(path))[0]
        image = _resize_nearest(_imread_rgb(path), self.resize_size)
        label = self.mapping[os.path.basename(os.path.dirname(path))]
        if self.fold == "train":
            image = self.augument(image)
        if self.raw_u8:
            tensor = torch.from_numpy(np.ascontiguousarray(image))            # uint8 HWC
        else:
            tensor = torch.from_numpy(self.normalize(image).transpose(2, 0, 1).copy()).float()
        return {"image": tensor, "label": label, "image_id": image_id}

    def augument(self, image: np.ndarray) -> np.ndarray:     # (sic) upstream spelling kept
        r = self._augment_rng()
        image = np.rot90(image, k=r.randrange(4))
        if r.random() > 0.5:
            image = image[::-1, ...]
        if r.random() > 0.5:
            image = image[:, ::-1, ...]
        f = lambda: 0.9 + r.random() * 0.2
        if r.random() > 0.95:
            image = saturation(image, f())
        elif r.random() > 0.95:
            image = brightness(image, f())
        elif r.random() > 0.95:
            image = contrast(image, f())
        return np.ascontiguousarray(image)

    @staticmethod
    def normalize(image, mean=IMAGENET_MEAN, std=IMAGENET_STD) -> np.ndarray:
        image = np.asarray(image, dtype="float32") / 255
        return (image - np.asarray(mean, dtype="float32")) / np.asarray(std, dtype="float32")


def make_loader(dataset: Dataset, batch_size: int, sampler: ShardedSampler, num_workers: int = 6,
                pin_memory: Optional[bool] = None) -> DataLoader:
    """DataLoader wired like train.py:114 (shuffle=False, sampler=..., pinned, 6 workers)."""
    if pin_memory is None:
        pin_memory = torch.cuda.is_available()
    return DataLoader(dataset, batch_size=batch_size, shuffle=False, num_workers=num_workers,
                      pin_memory=pin_memory, sampler=sampler, persistent_workers=num_workers > 0)


def write_fake_dataset(root: str, classes: int = 3, per_class: int = 4, size: int = 40,
                       folds=("train", "valid"), seed: int = 0) -> None:
    """Create a tiny PNG dataset in the reference's folder layout (used by tests / smoke runs)."""
    rng = np.random.default_rng(seed)
    for fold in folds:
        for c in range(classes):
            d = Path(root) / fold / f"class_{c}"
            d.mkdir(parents=True, exist_ok=True)
            for i in range(per_class):
                img = rng.integers(0, 256, (size, size, 3), dtype=np.uint8)
                img[..., c % 3] = np.minimum(255, img[..., c % 3].astype(int) + 80 * (c + 1) // classes)
                if cv2 is not None:
                    cv2.imwrite(str(d / f"img_{c}_{i}.png"), img[..., ::-1])
                else:
                    from PIL import Image
                    Image.fromarray(img).save(d / f"img_{c}_{i}.png")
