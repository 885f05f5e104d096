"""Reference-compatible module name: ``from utils import AverageMeter, Accuracy`` (reference utils.py:5-27)."""
from picd_b200.utils.meters import Accuracy, AverageMeter  # noqa: F401
