"""Reference-compatible import path: ``from dp.loader import ImageDataset`` (reference dp/loader.py:15)."""
from picd_b200.data.loader import ImageDataset, make_loader  # noqa: F401
from picd_b200.data.synthetic import SyntheticDeviceLoader, SyntheticHostLoader  # noqa: F401
from picd_b200.parallel.sampler import ShardedSampler  # noqa: F401
