"""Reference-compatible import path: ``from nn.classifier import Classifier`` (reference nn/classifier.py:7)."""
from picd_b200.models.classifier import BACKBONES, Classifier  # noqa: F401
