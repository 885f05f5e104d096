"""Data sources: the reference's folder dataset and the on-device synthetic generator."""
from .loader import (ImageDataset, brightness, contrast, make_loader, saturation, write_fake_dataset)
from .prefetch import DevicePrefetcher
from .synthetic import SyntheticDeviceLoader, SyntheticHostLoader

__all__ = [n for n in dir() if not n.startswith("_")]
