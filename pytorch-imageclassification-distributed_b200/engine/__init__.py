"""Training runtime: epoch loops and the command line."""
from .trainer import WeightedCrossEntropy, compute_loss, fit, train_epoch, val_epoch

__all__ = [n for n in dir() if not n.startswith("_")]
