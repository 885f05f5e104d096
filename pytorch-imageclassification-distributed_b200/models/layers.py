"""Parameter-holding building blocks.

Modules only own parameters/buffers (fp32 masters, torchvision-compatible names
so reference checkpoints with ``module.encoder.*`` keys load unchanged); the math
lives in :mod:`picd_b200.ops`, which fuses conv + BN + activation + residual.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
import torch.nn as nn

from .. import ops


# BatchNorm's ``num_batches_tracked`` counters: on CUDA the per-layer ``+= 1`` (one tiny kernel per BN layer per step, 53 for
# ResNet-50) is deferred and applied by ONE multi-tensor add at the end of the model's forward pass.
_pending_counters = []


def _count_batch(bn) -> None:
    t = bn.num_batches_tracked
    if t.is_cuda:
        _pending_counters.append(t)
    else:
        t += 1


def flush_batch_counters() -> None:
    """Apply the deferred ``num_batches_tracked += 1`` updates (called at the end of Classifier / backbone /
    FusedDataParallel forwards and before a BatchNorm state dict is read)."""
    if _pending_counters:
        torch._foreach_add_(_pending_counters, 1)
        _pending_counters.clear()


def _pair(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


class Conv2d(nn.Module):
    """Convolution weights [Cout, Cin/groups, R, S] kept channels_last, i.e. physically
    [Cout, R, S, Cin] -- the K-major operand layout the tcgen05 kernels read through TMA."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, groups=1, bias=False):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding = _pair(kernel_size), _pair(stride), _pair(padding)
        self.groups = groups
        w = torch.empty(out_channels, in_channels // groups, *self.kernel_size)
        nn.init.kaiming_normal_(w, mode="fan_out", nonlinearity="relu")
        self.weight = nn.Parameter(w.contiguous(memory_format=torch.channels_last))
        if bias:
            self.bias = nn.Parameter(torch.zeros(out_channels))
        else:
            self.register_parameter("bias", None)

    def forward(self, x):
        return ops.conv2d(x, self.weight, self.bias, self.stride, self.padding, self.groups)

    def extra_repr(self):
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, "
                f"stride={self.stride}, padding={self.padding}, groups={self.groups}")


class BatchNorm2d(nn.Module):
    """BatchNorm parameters + running statistics.  ``sync_group`` False = per-replica,
    True = WORLD, or a ProcessGroup (reference: nn.SyncBatchNorm.convert_sync_batchnorm,
    train.py:124)."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1):
        super().__init__()
        self.num_features, self.eps, self.momentum = num_features, eps, momentum
        self.weight = nn.Parameter(torch.ones(num_features))
        self.bias = nn.Parameter(torch.zeros(num_features))
        self.register_buffer("running_mean", torch.zeros(num_features))
        self.register_buffer("running_var", torch.ones(num_features))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        self.sync_group = False

    def forward(self, x, act=None, residual=None):
        if self.training:
            _count_batch(self)
        return ops.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var,
                              self.training, self.momentum, self.eps, self.sync_group, act, residual)

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        flush_batch_counters()
        return super()._save_to_state_dict(destination, prefix, keep_vars)

    def extra_repr(self):
        return f"{self.num_features}, eps={self.eps}, momentum={self.momentum}, sync={self.sync_group}"


class SyncBatchNorm(BatchNorm2d):
    """A BatchNorm2d whose batch statistics span the process group."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, process_group=None):
        super().__init__(num_features, eps, momentum)
        self.sync_group = True if process_group is None else process_group


def convert_sync_batchnorm(module: nn.Module, process_group=None) -> nn.Module:
    """In-place analogue of ``nn.SyncBatchNorm.convert_sync_batchnorm`` (train.py:124):
    every BatchNorm2d starts merging statistics across ``process_group`` (None = WORLD)."""
    for m in module.modules():
        if isinstance(m, BatchNorm2d):
            m.__class__ = SyncBatchNorm
            m.sync_group = True if process_group is None else process_group
    return module


def conv_bn_act(x, conv: Conv2d, bn: BatchNorm2d, act: Optional[str] = "relu", residual=None, grad_fork=None,
                padding=None):
    """Fused act(BN(conv(x)) + residual) using the parameters of a Conv2d/BatchNorm2d pair.  ``padding`` overrides the
    conv's own (used for TF-"same" layers: a (top, left, bottom, right) tuple)."""
    if bn.training:
        _count_batch(bn)
    return ops.conv_bn_act(
        x, conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var,
        stride=conv.stride, padding=conv.padding if padding is None else padding, groups=conv.groups,
        training=bn.training,
        momentum=bn.momentum, eps=bn.eps, act=act or "none", residual=residual,
        sync_group=bn.sync_group, grad_fork=grad_fork)


def conv_bn_relu_maxpool(x, conv: Conv2d, bn: BatchNorm2d):
    """Stem ``max_pool2d(relu(BN(conv(x))), 3, 2, 1)`` as one op (ops.conv_bn_relu_maxpool)."""
    if conv.groups != 1:
        return ops.max_pool2d(conv_bn_act(x, conv, bn, "relu"), 3, 2, 1)
    if bn.training:
        _count_batch(bn)
    return ops.conv_bn_relu_maxpool(x, conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var, stride=conv.stride,
                                    padding=conv.padding, training=bn.training, momentum=bn.momentum, eps=bn.eps,
                                    sync_group=bn.sync_group)


class Linear(nn.Module):
    def __init__(self, in_features, out_features, bias=True, act: Optional[str] = None):
        super().__init__()
        self.in_features, self.out_features, self.act = in_features, out_features, act
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if bias:
            bound = 1 / math.sqrt(in_features)
            self.bias = nn.Parameter(torch.empty(out_features).uniform_(-bound, bound))
        else:
            self.register_parameter("bias", None)

    def forward(self, x):
        return ops.linear(x, self.weight, self.bias, self.act)

    def extra_repr(self):
        return f"{self.in_features}, {self.out_features}, bias={self.bias is not None}, act={self.act}"


class ReLU(nn.Module):
    def __init__(self, inplace: bool = False):
        super().__init__()

    def forward(self, x):
        return ops.activation(x, "relu")


class LayerNorm(nn.Module):
    def __init__(self, dim, eps=1e-6):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim))
        self.bias = nn.Parameter(torch.zeros(dim))

    def forward(self, x):
        return ops.layer_norm(x, self.weight, self.bias, self.eps)


class Dropout(nn.Module):
    def __init__(self, p=0.5):
        super().__init__()
        self.p = p

    def forward(self, x):
        return ops.dropout(x, self.p, self.training)


class MLPHead(nn.Sequential):
    """The reference's classifier head: in -> 128 -> ReLU -> 64 -> ReLU -> 32 -> ReLU -> nc
    (nn/classifier.py:26-34).  Kept as a Sequential with Linear layers at indices 0,2,4,6 so
    state-dict keys (``fc.0.weight`` ...) match reference checkpoints.  The ReLUs are fused
    into the preceding Linear's epilogue."""

    def __init__(self, in_features: int, num_classes: int):
        super().__init__(
            Linear(in_features, 128, act="relu"), nn.Identity(),
            Linear(128, 64, act="relu"), nn.Identity(),
            Linear(64, 32, act="relu"), nn.Identity(),
            Linear(32, num_classes),
        )
        self.in_features = in_features


def prepare_input(x: torch.Tensor) -> torch.Tensor:
    """Bring a batch into the layout/dtype the backbone computes in: channels_last always,
    bf16 on CUDA (fp32 on CPU, where the torch oracle runs)."""
    if x.is_cuda and x.dtype != torch.bfloat16 and os.environ.get("PICD_REF_FP32", "0") != "1":
        x = x.to(torch.bfloat16)      # PICD_REF_FP32=1 (with PICD_FORCE_TORCH=1) keeps an fp32 oracle run in fp32
    if x.dim() == 4 and not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    return x
