"""``Classifier(name, num_classes)`` -- backbone + the reference's 4-layer MLP head.

Behaviour follows nn/classifier.py:7-37: the backbone lives in ``self.encoder``
(so checkpoints carry ``module.encoder.*`` keys), the final ``fc`` is replaced by
``in -> 128 -> 64 -> 32 -> num_classes`` with ReLUs (:26-34), inception also gets a
fresh ``AuxLogits.fc`` (:22-23), and ``forward`` is ``self.encoder(images)`` (:36-37).
Fixes carried from SURVEY.md 0.1: weights are random-init by default (D6, no
network), EfficientNet's head is ``_fc`` (D7), an unknown name raises a clear
error (D8); ``resnet18`` and ``vit_b_16`` (head swap on ``heads.head``) are added.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .efficientnet import efficientnet
from .inception import inception_v3
from .layers import Linear, MLPHead, flush_batch_counters
from .resnet import resnet18, resnet34, resnet50, resnet101
from .vit import vit_b_16, vit_tiny_test

BACKBONES = ("resnet18", "resnet34", "resnet50", "resnet101", "inceptionv3", "efficientnet-b0",
             "efficientnet-b1", "efficientnet-b2", "efficientnet-b3", "vit_b_16", "vit_tiny_test")


class Classifier(nn.Module):
    def __init__(self, name: str, num_classes: int, pretrained: Optional[str] = None,
                 image_size: Optional[int] = None, small_input: bool = False):
        super().__init__()
        self.name = name
        if num_classes < 1:
            raise ValueError("num_classes must be >= 1 (the reference's empty label mapping, "
                             "SURVEY.md D5, would build Linear(32, 0))")
        if name == "resnet18":
            self.encoder = resnet18(small_input=small_input)
        elif name == "resnet34":
            self.encoder = resnet34(small_input=small_input)
        elif name == "resnet50":
            self.encoder = resnet50(small_input=small_input)
        elif name == "resnet101":
            self.encoder = resnet101(small_input=small_input)
        elif name == "inceptionv3":
            # pretrained=True upstream forces transform_input=True
            self.encoder = inception_v3(transform_input=pretrained is not None)
            self.encoder.AuxLogits.fc = Linear(self.encoder.AuxLogits.fc.in_features, num_classes)
        elif name.startswith("efficientnet-b"):
            self.encoder = efficientnet(name)
        elif name == "vit_b_16":
            self.encoder = vit_b_16(image_size=image_size or 224)
        elif name == "vit_tiny_test":
            self.encoder = vit_tiny_test(image_size=image_size or 32)
        else:
            raise ValueError(f"unknown backbone {name!r}; choose one of {BACKBONES}")

        if name.startswith("vit"):
            self.encoder.heads.head = MLPHead(self.encoder.heads.head.in_features, num_classes)
        elif name.startswith("efficientnet"):
            self.encoder._fc = MLPHead(self.encoder._fc.in_features, num_classes)
        else:
            self.encoder.fc = MLPHead(self.encoder.fc.in_features, num_classes)

        if pretrained is not None:
            self.load_pretrained(pretrained)

    def load_pretrained(self, path: str) -> None:
        """Load backbone weights from a local torchvision-format state dict (offline
        replacement for ``pretrained=True``); head keys that do not match are skipped."""
        sd = torch.load(path, map_location="cpu")
        sd = sd.get("state_dict", sd)
        own = self.encoder.state_dict()
        keep = {k: v for k, v in sd.items() if k in own and own[k].shape == v.shape}
        self.encoder.load_state_dict(keep, strict=False)

    @property
    def returns_aux(self) -> bool:
        return self.name == "inceptionv3"

    def forward(self, images):
        out = self.encoder(images)
        flush_batch_counters()
        return out
