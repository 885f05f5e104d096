"""Pure-PyTorch restatements of every op in :mod:`picd_b200.ops`.

These serve three purposes: (1) the CPU / gloo execution path (BASELINE.json
config 1 runs ResNet-18 on two CPU ranks), (2) the numerical oracle each CUDA
kernel is tested against (tests compare a kernel with the fp32 evaluation of
the function of the same name here), and (3) the ``PICD_FORCE_TORCH=1`` A/B
switch on a GPU box.

Semantics follow the library calls the reference reaches through
``train.py:49-73``.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch
import torch.distributed as dist
import torch.nn.functional as F


# --------------------------------------------------------------------------- BN
class _SyncBatchNormRef(torch.autograd.Function):
    """Cross-replica batch norm usable on ANY backend (gloo/CPU included).

    torch's SyncBatchNorm refuses CPU modules under DDP
    (torch/nn/parallel/distributed.py:2259-2266), so the CPU plumbing config
    needs its own.  Statistics are merged by summing (sum, sumsq, count), which
    is equivalent to the count-weighted (mean, invstd, count) merge torch does
    in torch/nn/modules/_functions.py:59-115 and is what the CUDA kernel fuses
    over NVLink peer memory.
    """

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, momentum, eps, group):
        c = x.shape[1]
        xf = x.float()
        dims = [0] + list(range(2, x.dim()))
        count = torch.tensor([xf.numel() / c], dtype=torch.float32, device=x.device)
        packed = torch.cat([xf.sum(dims), (xf * xf).sum(dims), count])
        dist.all_reduce(packed, group=group)
        n = packed[-1]
        mean = packed[:c] / n
        var = (packed[c:2 * c] / n - mean * mean).clamp_min_(0.0)
        invstd = torch.rsqrt(var + eps)
        if running_mean is not None:
            with torch.no_grad():
                unbiased = var * (n / (n - 1).clamp_min(1.0))
                running_mean.mul_(1 - momentum).add_(mean.to(running_mean.dtype), alpha=momentum)
                running_var.mul_(1 - momentum).add_(unbiased.to(running_var.dtype), alpha=momentum)
        shape = [1, c] + [1] * (x.dim() - 2)
        xhat = (xf - mean.view(shape)) * invstd.view(shape)
        out = xhat * gamma.float().view(shape) + beta.float().view(shape)
        ctx.save_for_backward(xhat, gamma, invstd, n)
        ctx.group = group
        return out.to(x.dtype)

    @staticmethod
    def backward(ctx, dout):
        xhat, gamma, invstd, n = ctx.saved_tensors
        c = xhat.shape[1]
        dims = [0] + list(range(2, xhat.dim()))
        g = dout.float()
        sum_dy = g.sum(dims)
        sum_dy_xhat = (g * xhat).sum(dims)
        # parameter gradients use the LOCAL sums (DDP averages them afterwards),
        # input gradients use the GLOBAL ones (torch/nn/modules/_functions.py:144-179)
        dgamma, dbeta = sum_dy_xhat.clone(), sum_dy.clone()
        packed = torch.cat([sum_dy, sum_dy_xhat])
        dist.all_reduce(packed, group=ctx.group)
        shape = [1, c] + [1] * (xhat.dim() - 2)
        mean_dy = (packed[:c] / n).view(shape)
        mean_dy_xhat = (packed[c:] / n).view(shape)
        dx = (g - mean_dy - xhat * mean_dy_xhat) * (gamma.float() * invstd).view(shape)
        return dx.to(dout.dtype), dgamma.to(gamma.dtype), dbeta.to(gamma.dtype), None, None, None, None, None


def _group_size(group) -> int:
    if group is False or not (dist.is_available() and dist.is_initialized()):
        return 1
    return dist.get_world_size(group if group is not True else None)


def batch_norm(x, gamma, beta, running_mean, running_var, training: bool, momentum: float,
               eps: float, sync_group=False):
    """BatchNorm over dim 1.  ``sync_group``: False = local, True = WORLD, or a ProcessGroup."""
    if training and _group_size(sync_group) > 1:
        grp = None if sync_group is True else sync_group
        return _SyncBatchNormRef.apply(x, gamma, beta, running_mean, running_var, momentum, eps, grp)
    return F.batch_norm(x, running_mean, running_var, gamma, beta, training, momentum, eps)


def conv2d(x, w, bias=None, stride=1, padding=0, groups: int = 1):
    return F.conv2d(x, w.to(x.dtype), None if bias is None else bias.to(x.dtype), stride, padding, 1, groups)


def conv_bn_act(x, w, gamma, beta, running_mean, running_var, *, stride=1, padding=0, groups=1,
                training=True, momentum=0.1, eps=1e-5, act: str = "relu", residual=None,
                sync_group=False, conv_bias=None):
    """act(BN(conv(x, w)) + residual) -- the fused unit of every CNN backbone here."""
    y = conv2d(x, w, conv_bias, stride, padding, groups)
    z = batch_norm(y, gamma, beta, running_mean, running_var, training, momentum, eps, sync_group)
    if residual is not None:
        z = z + residual
    return activation(z, act)


def activation(x, act: Optional[str]):
    if act is None or act == "none":
        return x
    if act == "relu":
        return F.relu(x)
    if act == "gelu":
        return F.gelu(x)
    if act in ("swish", "silu"):
        return F.silu(x)
    if act == "sigmoid":
        return torch.sigmoid(x)
    raise ValueError(f"unknown activation {act!r}")


def linear(x, w, b=None, act: Optional[str] = None):
    y = F.linear(x, w.to(x.dtype), None if b is None else b.to(x.dtype))
    return activation(y, act)


# ------------------------------------------------------------------------ pools
def max_pool2d(x, kernel_size, stride, padding=0, ceil_mode=False):
    return F.max_pool2d(x, kernel_size, stride, padding, ceil_mode=ceil_mode)


def avg_pool2d(x, kernel_size, stride, padding=0, count_include_pad=True):
    if x.is_cuda and x.dim() == 4 and not x.is_contiguous():
        # torch 2.11's CUDA avg_pool2d BACKWARD for channels_last inputs ignores the padding offset (measured on B200: a
        # unit gradient at output (0, 0) of a 3x3/s1/p1 pool lands on inputs [0..2, 0..2] instead of [0..1, 0..1]; the
        # NCHW kernel and picd_pool_bwd agree with the definition).  The oracle therefore pools in NCHW.
        y = F.avg_pool2d(x.contiguous(), kernel_size, stride, padding, count_include_pad=count_include_pad)
        return y.contiguous(memory_format=torch.channels_last)
    return F.avg_pool2d(x, kernel_size, stride, padding, count_include_pad=count_include_pad)


def global_avg_pool(x):
    """[N,C,H,W] -> [N,C] (mean over H,W, fp32 accumulate)."""
    return x.float().mean(dim=(2, 3)).to(x.dtype)


# ------------------------------------------------------------------------- loss
def cross_entropy(logits, labels, weight=None):
    """Class-weighted softmax cross-entropy, mean = sum(w_i l_i) / sum(w_i)
    (reference train.py:157-158)."""
    return F.cross_entropy(logits.float(), labels, None if weight is None else weight.float())


# -------------------------------------------------------------------------- ViT
def layer_norm(x, gamma, beta, eps: float = 1e-6):
    return F.layer_norm(x.float(), (x.shape[-1],), gamma.float(), beta.float(), eps).to(x.dtype)


def attention(q, k, v, scale: Optional[float] = None):
    """softmax(q k^T * scale) v for [B, heads, S, D] tensors (no mask, no dropout)."""
    d = q.shape[-1]
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    s = torch.matmul(q.float(), k.float().transpose(-1, -2)) * scale
    p = torch.softmax(s, dim=-1)
    return torch.matmul(p, v.float()).to(q.dtype)


def dropout(x, p: float, training: bool):
    return F.dropout(x, p, training)


# ------------------------------------------------------------------- optimizers
def sgd_momentum_step(p, g, buf, lr, momentum, weight_decay, nesterov=False, grad_scale=1.0):
    """torch.optim.SGD semantics (dampening 0): g' = g*scale + wd*p; buf = mu*buf + g'; p -= lr*buf."""
    g = g.float() * grad_scale
    if weight_decay:
        g = g + weight_decay * p
    buf.mul_(momentum).add_(g)
    upd = g + momentum * buf if nesterov else buf
    p.add_(upd, alpha=-lr)


def adam_step(p, g, m, v, step: int, lr, beta1, beta2, eps, weight_decay, decoupled: bool, grad_scale=1.0):
    """torch.optim.Adam (decoupled=False, L2 into the gradient) / AdamW (decoupled=True)."""
    g = g.float() * grad_scale
    if weight_decay:
        if decoupled:
            p.mul_(1 - lr * weight_decay)
        else:
            g = g + weight_decay * p
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


# --------------------------------------------------------------- synthetic data
IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def _hash_u32(x: torch.Tensor) -> torch.Tensor:
    """The 32-bit integer hash the device generator uses (lowbias32), on int64 lanes."""
    m = 0xFFFFFFFF
    x = x & m
    x = (x ^ (x >> 16)) & m
    x = (x * 0x7FEB352D) & m
    x = (x ^ (x >> 15)) & m
    x = (x * 0x846CA68B) & m
    x = (x ^ (x >> 16)) & m
    return x


def synthetic_batch(batch: int, channels: int, height: int, width: int, num_classes: int, seed: int,
                    step: int, device="cpu", dtype=torch.float32) -> Tuple[torch.Tensor, torch.Tensor]:
    """Deterministic pseudo-random uint8-like pixels, ImageNet-normalised (dp/loader.py:86-91).

    pixel(n,c,h,w) = hash(seed, step, linear NHWC index) & 255; label(n) = hash(...) % classes.
    The CUDA generator produces bit-identical values directly in NHWC bf16.
    """
    n_el = batch * height * width * channels
    idx = torch.arange(n_el, dtype=torch.int64, device=device)
    base = (seed * 0x9E3779B1 + step * 0x85EBCA77) & 0xFFFFFFFF
    px = (_hash_u32(idx + base) & 255).to(torch.float32) / 255.0
    img = px.view(batch, height, width, channels)
    mean = torch.tensor(IMAGENET_MEAN[:channels], device=device)
    std = torch.tensor(IMAGENET_STD[:channels], device=device)
    img = ((img - mean) / std).permute(0, 3, 1, 2)
    lab_idx = torch.arange(batch, dtype=torch.int64, device=device)
    labels = _hash_u32(lab_idx + ((base ^ 0xA5A5A5A5) & 0xFFFFFFFF)) % num_classes
    return img.to(dtype).contiguous(memory_format=torch.channels_last), labels


def normalize_u8(images_u8: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    """uint8 NHWC -> normalised NCHW-logical channels_last tensor (dp/loader.py:86-91)."""
    c = images_u8.shape[-1]
    mean = torch.tensor(IMAGENET_MEAN[:c], device=images_u8.device)
    std = torch.tensor(IMAGENET_STD[:c], device=images_u8.device)
    x = (images_u8.float() / 255.0 - mean) / std
    return x.permute(0, 3, 1, 2).to(dtype).contiguous(memory_format=torch.channels_last)
