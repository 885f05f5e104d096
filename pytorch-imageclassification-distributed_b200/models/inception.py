"""Inception-v3 (the reference's default backbone, train.py:122 / nn/classifier.py:20-23).

Same topology and state-dict keys as torchvision's ``inception_v3`` (96 conv+BN
pairs, aux head on the 768-channel 17x17 map), but each BasicConv2d is the fused
conv+BN+ReLU op and every asymmetric (1x7 / 7x1 / 1x3 / 3x1 / 5x5) filter goes
through the same tap-list implicit-GEMM kernel.  In train mode ``forward`` returns
``(logits, aux_logits)`` as the reference expects (train.py:48-52).
"""
from __future__ import annotations

from collections import namedtuple

import torch
import torch.nn as nn

from .. import ops
from .layers import BatchNorm2d, Conv2d, Dropout, Linear, conv_bn_act, prepare_input

InceptionOutputs = namedtuple("InceptionOutputs", ["logits", "aux_logits"])


class BasicConv2d(nn.Module):
    def __init__(self, cin, cout, kernel_size, stride=1, padding=0, std=0.1):
        super().__init__()
        self.conv = Conv2d(cin, cout, kernel_size, stride, padding)
        nn.init.trunc_normal_(self.conv.weight, mean=0.0, std=std, a=-2, b=2)
        self.bn = BatchNorm2d(cout, eps=0.001)

    def forward(self, x, grad_fork=None):
        return conv_bn_act(x, self.conv, self.bn, "relu", grad_fork=grad_fork)


def _acc(fk):
    """grad_fork argument of a branch whose input gradient is ADDED into the one deposited by the block's last branch."""
    return None if fk is None else (fk, "accumulate")


def _cat(xs):
    return ops.concat_channels(xs)


class InceptionA(nn.Module):
    def __init__(self, cin, pool_features):
        super().__init__()
        self.branch1x1 = BasicConv2d(cin, 64, 1)
        self.branch5x5_1 = BasicConv2d(cin, 48, 1)
        self.branch5x5_2 = BasicConv2d(48, 64, 5, padding=2)
        self.branch3x3dbl_1 = BasicConv2d(cin, 64, 1)
        self.branch3x3dbl_2 = BasicConv2d(64, 96, 3, padding=1)
        self.branch3x3dbl_3 = BasicConv2d(96, 96, 3, padding=1)
        self.branch_pool = BasicConv2d(cin, pool_features, 1)

    def forward(self, x):
        # the block input feeds four branches: the pooling branch (created last, so its backward runs first) deposits its
        # input gradient, the three 1x1 convs add theirs into it with TMA reduce-add stores (ops.fork_n)
        (x1, x2, x3, x4), fk = ops.fork_n(x, 4)
        b1 = self.branch1x1(x1, _acc(fk))
        b5 = self.branch5x5_2(self.branch5x5_1(x2, _acc(fk)))
        b3 = self.branch3x3dbl_3(self.branch3x3dbl_2(self.branch3x3dbl_1(x3, _acc(fk))))
        bp = self.branch_pool(ops.avg_pool2d(x4, 3, 1, 1, grad_fork=fk))
        return _cat([b1, b5, b3, bp])


class InceptionB(nn.Module):
    def __init__(self, cin):
        super().__init__()
        self.branch3x3 = BasicConv2d(cin, 384, 3, stride=2)
        self.branch3x3dbl_1 = BasicConv2d(cin, 64, 1)
        self.branch3x3dbl_2 = BasicConv2d(64, 96, 3, padding=1)
        self.branch3x3dbl_3 = BasicConv2d(96, 96, 3, stride=2)

    def forward(self, x):
        (x1, x2, x3), fk = ops.fork_n(x, 3)
        b3 = self.branch3x3(x1, _acc(fk))
        bd = self.branch3x3dbl_3(self.branch3x3dbl_2(self.branch3x3dbl_1(x2, _acc(fk))))
        return _cat([b3, bd, ops.max_pool2d(x3, 3, 2, grad_fork=fk)])


class InceptionC(nn.Module):
    def __init__(self, cin, c7):
        super().__init__()
        self.branch1x1 = BasicConv2d(cin, 192, 1)
        self.branch7x7_1 = BasicConv2d(cin, c7, 1)
        self.branch7x7_2 = BasicConv2d(c7, c7, (1, 7), padding=(0, 3))
        self.branch7x7_3 = BasicConv2d(c7, 192, (7, 1), padding=(3, 0))
        self.branch7x7dbl_1 = BasicConv2d(cin, c7, 1)
        self.branch7x7dbl_2 = BasicConv2d(c7, c7, (7, 1), padding=(3, 0))
        self.branch7x7dbl_3 = BasicConv2d(c7, c7, (1, 7), padding=(0, 3))
        self.branch7x7dbl_4 = BasicConv2d(c7, c7, (7, 1), padding=(3, 0))
        self.branch7x7dbl_5 = BasicConv2d(c7, 192, (1, 7), padding=(0, 3))
        self.branch_pool = BasicConv2d(cin, 192, 1)

    def forward(self, x):
        (x1, x2, x3, x4), fk = ops.fork_n(x, 4)
        b1 = self.branch1x1(x1, _acc(fk))
        b7 = self.branch7x7_3(self.branch7x7_2(self.branch7x7_1(x2, _acc(fk))))
        bd = self.branch7x7dbl_5(self.branch7x7dbl_4(self.branch7x7dbl_3(
            self.branch7x7dbl_2(self.branch7x7dbl_1(x3, _acc(fk))))))
        bp = self.branch_pool(ops.avg_pool2d(x4, 3, 1, 1, grad_fork=fk))
        return _cat([b1, b7, bd, bp])


class InceptionD(nn.Module):
    def __init__(self, cin):
        super().__init__()
        self.branch3x3_1 = BasicConv2d(cin, 192, 1)
        self.branch3x3_2 = BasicConv2d(192, 320, 3, stride=2)
        self.branch7x7x3_1 = BasicConv2d(cin, 192, 1)
        self.branch7x7x3_2 = BasicConv2d(192, 192, (1, 7), padding=(0, 3))
        self.branch7x7x3_3 = BasicConv2d(192, 192, (7, 1), padding=(3, 0))
        self.branch7x7x3_4 = BasicConv2d(192, 192, 3, stride=2)

    def forward(self, x):
        (x1, x2, x3), fk = ops.fork_n(x, 3)
        b3 = self.branch3x3_2(self.branch3x3_1(x1, _acc(fk)))
        b7 = self.branch7x7x3_4(self.branch7x7x3_3(self.branch7x7x3_2(self.branch7x7x3_1(x2, _acc(fk)))))
        return _cat([b3, b7, ops.max_pool2d(x3, 3, 2, grad_fork=fk)])


class InceptionE(nn.Module):
    def __init__(self, cin):
        super().__init__()
        self.branch1x1 = BasicConv2d(cin, 320, 1)
        self.branch3x3_1 = BasicConv2d(cin, 384, 1)
        self.branch3x3_2a = BasicConv2d(384, 384, (1, 3), padding=(0, 1))
        self.branch3x3_2b = BasicConv2d(384, 384, (3, 1), padding=(1, 0))
        self.branch3x3dbl_1 = BasicConv2d(cin, 448, 1)
        self.branch3x3dbl_2 = BasicConv2d(448, 384, 3, padding=1)
        self.branch3x3dbl_3a = BasicConv2d(384, 384, (1, 3), padding=(0, 1))
        self.branch3x3dbl_3b = BasicConv2d(384, 384, (3, 1), padding=(1, 0))
        self.branch_pool = BasicConv2d(cin, 192, 1)

    def forward(self, x):
        (x1, x2, x3, x4), fk = ops.fork_n(x, 4)
        b1 = self.branch1x1(x1, _acc(fk))
        b3 = self.branch3x3_1(x2, _acc(fk))
        (b3a, b3b), fk3 = ops.fork_n(b3, 2)        # inner two-way forks: the second conv deposits, the first accumulates
        b3 = _cat([self.branch3x3_2a(b3a, _acc(fk3)), self.branch3x3_2b(b3b, None if fk3 is None else (fk3, "deposit_dx"))])
        bd = self.branch3x3dbl_2(self.branch3x3dbl_1(x3, _acc(fk)))
        (bda, bdb), fkd = ops.fork_n(bd, 2)
        bd = _cat([self.branch3x3dbl_3a(bda, _acc(fkd)), self.branch3x3dbl_3b(bdb, None if fkd is None else (fkd, "deposit_dx"))])
        bp = self.branch_pool(ops.avg_pool2d(x4, 3, 1, 1, grad_fork=fk))
        return _cat([b1, b3, bd, bp])


class InceptionAux(nn.Module):
    def __init__(self, cin, num_classes):
        super().__init__()
        self.conv0 = BasicConv2d(cin, 128, 1)
        self.conv1 = BasicConv2d(128, 768, 5, std=0.01)
        self.fc = Linear(768, num_classes)
        nn.init.trunc_normal_(self.fc.weight, mean=0.0, std=0.001, a=-2, b=2)

    def forward(self, x):
        x = ops.avg_pool2d(x, 5, 3)
        x = self.conv1(self.conv0(x))
        return self.fc(ops.global_avg_pool(x))


class Inception3(nn.Module):
    def __init__(self, num_classes=1000, aux_logits=True, transform_input=False, dropout=0.5):
        super().__init__()
        self.aux_logits, self.transform_input = aux_logits, transform_input
        self.register_buffer("_ti_scale", torch.tensor([0.229 / 0.5, 0.224 / 0.5, 0.225 / 0.5]).view(1, 3, 1, 1),
                             persistent=False)
        self.register_buffer("_ti_shift", torch.tensor([(0.485 - 0.5) / 0.5, (0.456 - 0.5) / 0.5,
                                                        (0.406 - 0.5) / 0.5]).view(1, 3, 1, 1), persistent=False)
        self.Conv2d_1a_3x3 = BasicConv2d(3, 32, 3, stride=2)
        self.Conv2d_2a_3x3 = BasicConv2d(32, 32, 3)
        self.Conv2d_2b_3x3 = BasicConv2d(32, 64, 3, padding=1)
        self.Conv2d_3b_1x1 = BasicConv2d(64, 80, 1)
        self.Conv2d_4a_3x3 = BasicConv2d(80, 192, 3)
        self.Mixed_5b = InceptionA(192, 32)
        self.Mixed_5c = InceptionA(256, 64)
        self.Mixed_5d = InceptionA(288, 64)
        self.Mixed_6a = InceptionB(288)
        self.Mixed_6b = InceptionC(768, 128)
        self.Mixed_6c = InceptionC(768, 160)
        self.Mixed_6d = InceptionC(768, 160)
        self.Mixed_6e = InceptionC(768, 192)
        self.AuxLogits = InceptionAux(768, num_classes) if aux_logits else None
        self.Mixed_7a = InceptionD(768)
        self.Mixed_7b = InceptionE(1280)
        self.Mixed_7c = InceptionE(2048)
        self.dropout = Dropout(dropout)
        self.fc = Linear(2048, num_classes)
        nn.init.trunc_normal_(self.fc.weight, mean=0.0, std=0.1, a=-2, b=2)

    def _transform_input(self, x):
        # torchvision applies this when pretrained weights are requested:
        # re-normalise from ImageNet mean/std to the (0.5, 0.5) convention per channel.
        # the constants live in (non-persistent) buffers: no host -> device copy per step, CUDA-graph safe
        return x * self._ti_scale.to(x.dtype) + self._ti_shift.to(x.dtype)

    def forward(self, x):
        if self.transform_input:
            x = self._transform_input(x)
        x = prepare_input(x)
        x = self.Conv2d_2b_3x3(self.Conv2d_2a_3x3(self.Conv2d_1a_3x3(x)))
        x = ops.max_pool2d(x, 3, 2)
        x = self.Conv2d_4a_3x3(self.Conv2d_3b_1x1(x))
        x = ops.max_pool2d(x, 3, 2)
        x = self.Mixed_5d(self.Mixed_5c(self.Mixed_5b(x)))
        x = self.Mixed_6a(x)
        x = self.Mixed_6e(self.Mixed_6d(self.Mixed_6c(self.Mixed_6b(x))))
        aux = None
        if self.AuxLogits is not None and self.training:
            aux = self.AuxLogits(x)
        x = self.Mixed_7c(self.Mixed_7b(self.Mixed_7a(x)))
        x = self.dropout(ops.global_avg_pool(x))
        x = self.fc(x)
        if self.training and self.aux_logits:
            return InceptionOutputs(x, aux)
        return x


def inception_v3(num_classes=1000, aux_logits=True, transform_input=False, **kw):
    return Inception3(num_classes, aux_logits, transform_input, **kw)
