"""EfficientNet (B0-B3) with ``efficientnet_pytorch`` parameter names.

The reference's ``efficientnet-b3`` branch calls ``EfficientNet.from_pretrained``
(nn/classifier.py:17-18) and then trips over that package naming its head ``_fc``
(SURVEY.md defect D7).  This restates the architecture (MBConv + squeeze-excite,
TF-style "same" padding, swish) with the same state-dict keys (``_conv_stem``,
``_blocks.N._expand_conv`` ...) so such checkpoints load, and exposes the head as
``_fc`` for :class:`picd_b200.models.Classifier` to swap.

Depthwise convolutions have no GEMM shape (one input channel per output channel):
they run on a native bandwidth-bound stencil kernel (csrc/ops/depthwise.cu) that also
takes the asymmetric TF-"same" padding directly; the pointwise convs, BN, swish, SE and
head go through the same fused ops as the other backbones.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from .layers import BatchNorm2d, Conv2d, Dropout, Linear, conv_bn_act, prepare_input

# (repeats, kernel, stride, expand, in, out, se_ratio)
_BASE_BLOCKS = [
    (1, 3, 1, 1, 32, 16, 0.25), (2, 3, 2, 6, 16, 24, 0.25), (2, 5, 2, 6, 24, 40, 0.25),
    (3, 3, 2, 6, 40, 80, 0.25), (3, 5, 1, 6, 80, 112, 0.25), (4, 5, 2, 6, 112, 192, 0.25),
    (1, 3, 1, 6, 192, 320, 0.25),
]
# name -> (width, depth, resolution, dropout)
_PARAMS = {
    "efficientnet-b0": (1.0, 1.0, 224, 0.2), "efficientnet-b1": (1.0, 1.1, 240, 0.2),
    "efficientnet-b2": (1.1, 1.2, 260, 0.3), "efficientnet-b3": (1.2, 1.4, 300, 0.3),
}


def _round_filters(f, width, divisor=8):
    f *= width
    new_f = max(divisor, int(f + divisor / 2) // divisor * divisor)
    if new_f < 0.9 * f:
        new_f += divisor
    return int(new_f)


def _same_pad_amounts(x, k, s):
    """(top, left, bottom, right) of TF 'SAME' padding for the current spatial size."""
    ih, iw = x.shape[-2:]
    ph = max((math.ceil(ih / s) - 1) * s + k - ih, 0)
    pw = max((math.ceil(iw / s) - 1) * s + k - iw, 0)
    return (ph // 2, pw // 2, ph - ph // 2, pw - pw // 2)


def _same_pad(x, k, s):
    """TF 'SAME' padding (possibly asymmetric) applied explicitly; returns padded x."""
    ih, iw = x.shape[-2:]
    ph = max((math.ceil(ih / s) - 1) * s + k - ih, 0)
    pw = max((math.ceil(iw / s) - 1) * s + k - iw, 0)
    if ph or pw:
        x = F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2])
        if x.dim() == 4:
            x = x.contiguous(memory_format=torch.channels_last)
    return x


class MBConv(nn.Module):
    def __init__(self, cin, cout, k, s, expand, se_ratio, drop_connect):
        super().__init__()
        self.k, self.s, self.expand, self.drop_connect = k, s, expand, drop_connect
        self.skip = (s == 1 and cin == cout)
        mid = cin * expand
        if expand != 1:
            self._expand_conv = Conv2d(cin, mid, 1)
            self._bn0 = BatchNorm2d(mid, eps=1e-3, momentum=0.01)
        self._depthwise_conv = Conv2d(mid, mid, k, s, 0, groups=mid)
        self._bn1 = BatchNorm2d(mid, eps=1e-3, momentum=0.01)
        sq = max(1, int(cin * se_ratio))
        self._se_reduce = Conv2d(mid, sq, 1, bias=True)
        self._se_expand = Conv2d(sq, mid, 1, bias=True)
        self._project_conv = Conv2d(mid, cout, 1)
        self._bn2 = BatchNorm2d(cout, eps=1e-3, momentum=0.01)

    def forward(self, x):
        inp = x
        if self.expand != 1:
            x = conv_bn_act(x, self._expand_conv, self._bn0, "swish")
        x = conv_bn_act(x, self._depthwise_conv, self._bn1, "swish", padding=_same_pad_amounts(x, self.k, self.s))
        # squeeze-excite on the pooled [N, C] vector: two tiny GEMMs
        s = ops.global_avg_pool(x)
        s = ops.linear(s, self._se_reduce.weight.flatten(1), self._se_reduce.bias, "swish")
        s = ops.linear(s, self._se_expand.weight.flatten(1), self._se_expand.bias, "sigmoid")
        residual, mask = None, None
        if self.skip:
            residual = inp
            if self.training and self.drop_connect > 0:   # stochastic depth on the branch: folded into the gate kernel
                keep = 1.0 - self.drop_connect
                mask = (torch.rand(x.shape[0], device=x.device) < keep).float() / keep
        x = ops.channel_scale(x, s, mask)
        return conv_bn_act(x, self._project_conv, self._bn2, None, residual=residual)


class EfficientNet(nn.Module):
    def __init__(self, name="efficientnet-b3", num_classes=1000, drop_connect_rate=0.2):
        super().__init__()
        width, depth, self.resolution, dropout = _PARAMS[name]
        stem = _round_filters(32, width)
        self._conv_stem = Conv2d(3, stem, 3, 2, 0)
        self._bn0 = BatchNorm2d(stem, eps=1e-3, momentum=0.01)
        blocks = []
        total = sum(int(math.ceil(depth * r)) for r, *_ in _BASE_BLOCKS)
        for r, k, s, e, ci, co, se in _BASE_BLOCKS:
            ci, co = _round_filters(ci, width), _round_filters(co, width)
            for i in range(int(math.ceil(depth * r))):
                dc = drop_connect_rate * len(blocks) / total
                blocks.append(MBConv(ci if i == 0 else co, co, k, s if i == 0 else 1, e, se, dc))
        self._blocks = nn.ModuleList(blocks)
        head = _round_filters(1280, width)
        self._conv_head = Conv2d(_round_filters(320, width), head, 1)
        self._bn1 = BatchNorm2d(head, eps=1e-3, momentum=0.01)
        self._dropout = Dropout(dropout)
        self._fc = Linear(head, num_classes)

    def forward(self, x):
        x = prepare_input(x)
        x = conv_bn_act(_same_pad(x, 3, 2), self._conv_stem, self._bn0, "swish")
        for blk in self._blocks:
            x = blk(x)
        x = conv_bn_act(x, self._conv_head, self._bn1, "swish")
        x = self._dropout(ops.global_avg_pool(x))
        return self._fc(x)


def efficientnet(name="efficientnet-b3", num_classes=1000, **kw):
    return EfficientNet(name, num_classes, **kw)
