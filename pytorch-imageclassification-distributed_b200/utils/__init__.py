"""Meters, checkpointing, tracing and clock sampling."""
from .checkpoint import (BEST, LATEST, checkpoint_path, load_checkpoint, prefixed_state_dict,
                         save_checkpoint)
from .meters import Accuracy, AverageMeter

__all__ = [n for n in dir() if not n.startswith("_")]
from .lagged import LaggedScalarReader  # noqa: E402,F401
