"""ResNet-18/34/50/101 with torchvision-compatible parameter names.

The reference builds ``models.resnet50/resnet101`` (nn/classifier.py:11-15); the
driver configs add ResNet-18.  Structure (v1.5: stride on the 3x3 conv) and
state-dict keys match torchvision so ``module.encoder.*`` checkpoints load, but
every conv+BN(+ReLU)(+residual) runs as one fused op.
"""
from __future__ import annotations

from typing import List, Optional, Type

import torch
import torch.nn as nn

from .. import ops
from .layers import BatchNorm2d, Conv2d, Linear, conv_bn_act, conv_bn_relu_maxpool, prepare_input


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = Conv2d(inplanes, planes, 3, stride, 1)
        self.bn1 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, 3, 1, 1)
        self.bn2 = BatchNorm2d(planes)
        self.downsample = downsample

    def forward(self, x):
        # the block input feeds two branches; ops.fork lets the first conv's dgrad add straight into the skip gradient
        x1, x2, fk = ops.fork(x)
        out = conv_bn_act(x1, self.conv1, self.bn1, "relu", grad_fork=(fk, "accumulate"))
        if self.downsample is not None:      # created after conv1 so that its backward (the deposit) runs first
            identity = conv_bn_act(x2, self.downsample[0], self.downsample[1], act=None, grad_fork=(fk, "deposit_dx"))
            return conv_bn_act(out, self.conv2, self.bn2, "relu", residual=identity)
        return conv_bn_act(out, self.conv2, self.bn2, "relu", residual=x2, grad_fork=(fk, "deposit_res"))


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = Conv2d(inplanes, planes, 1)
        self.bn1 = BatchNorm2d(planes)
        self.conv2 = Conv2d(planes, planes, 3, stride, 1)
        self.bn2 = BatchNorm2d(planes)
        self.conv3 = Conv2d(planes, planes * 4, 1)
        self.bn3 = BatchNorm2d(planes * 4)
        self.downsample = downsample

    def forward(self, x):
        x1, x2, fk = ops.fork(x)
        out = conv_bn_act(x1, self.conv1, self.bn1, "relu", grad_fork=(fk, "accumulate"))
        out = conv_bn_act(out, self.conv2, self.bn2, "relu")
        if self.downsample is not None:      # created after conv1/conv2 so that its backward (the deposit) runs first
            identity = conv_bn_act(x2, self.downsample[0], self.downsample[1], act=None, grad_fork=(fk, "deposit_dx"))
            return conv_bn_act(out, self.conv3, self.bn3, "relu", residual=identity)
        return conv_bn_act(out, self.conv3, self.bn3, "relu", residual=x2, grad_fork=(fk, "deposit_res"))


class ResNet(nn.Module):
    def __init__(self, block: Type[nn.Module], layers: List[int], num_classes=1000, small_input=False):
        super().__init__()
        self.inplanes = 64
        self.small_input = small_input
        if small_input:  # CIFAR-shaped 32x32 batches (BASELINE.json config 1): 3x3 stem, no max-pool
            self.conv1 = Conv2d(3, 64, 3, 1, 1)
        else:
            self.conv1 = Conv2d(3, 64, 7, 2, 3)
        self.bn1 = BatchNorm2d(64)
        self.layer1 = self._make_layer(block, 64, layers[0], 1)
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.fc = Linear(512 * block.expansion, num_classes)
        for m in self.modules():  # zero-init is NOT applied (torchvision default zero_init_residual=False)
            if isinstance(m, BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)

    def _make_layer(self, block, planes, blocks, stride):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(Conv2d(self.inplanes, planes * block.expansion, 1, stride),
                                       BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        layers += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*layers)

    def forward(self, x):
        x = prepare_input(x)
        if self.small_input:
            x = conv_bn_act(x, self.conv1, self.bn1, "relu")
        else:
            x = conv_bn_relu_maxpool(x, self.conv1, self.bn1)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = ops.global_avg_pool(x)
        return self.fc(x)


def resnet18(num_classes=1000, **kw):
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes, **kw)


def resnet34(num_classes=1000, **kw):
    return ResNet(BasicBlock, [3, 4, 6, 3], num_classes, **kw)


def resnet50(num_classes=1000, **kw):
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes, **kw)


def resnet101(num_classes=1000, **kw):
    return ResNet(Bottleneck, [3, 4, 23, 3], num_classes, **kw)
