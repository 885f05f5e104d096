"""Operator front-end: every op dispatches to the sm_100a kernel library for CUDA
tensors and to the pure-torch oracle (:mod:`.reference`) otherwise.

There is exactly one device code path (no multi-backend dispatch): a CUDA tensor
runs the hand-written kernel.  The few shapes no model of this repository reaches and no
kernel covers (generic grouped convolutions, ``ceil_mode`` max-pooling, average pooling
that excludes the padding, attention with a head size other than 64 or more than 256
tokens) run the library operator and say so ONCE on stderr (``PICD_STRICT=1`` turns
that into an error) -- never silently.  ``PICD_FORCE_TORCH=1`` is an explicit debugging
switch that runs the oracle on the GPU for A/B comparison.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import reference as ref
from ._lib import lib_available, lib_path, require_lib  # noqa: F401


def _force_torch() -> bool:
    return os.environ.get("PICD_FORCE_TORCH", "0") == "1"


_warned = set()


def _library_fallback(what: str) -> None:
    """A CUDA tensor is about to run a library (ATen / cuDNN) operator because no sm_100a kernel covers this shape."""
    if os.environ.get("PICD_STRICT", "0") == "1":
        raise NotImplementedError(f"picd_b200: no native kernel for {what} (PICD_STRICT=1)")
    if what not in _warned:
        _warned.add(what)
        import sys
        print(f"picd_b200: WARNING: {what} has no sm_100a kernel here; running the library operator", file=sys.stderr)


def use_native(x: torch.Tensor) -> bool:
    """True when ``x`` must go through libpicd_b200.so (fails loudly if it is absent)."""
    if not x.is_cuda or _force_torch():
        return False
    require_lib()
    return True


# ---------------------------------------------------------------------- CNN ops
def fork(x):
    """Split an activation that feeds two branches (a residual block's input) into two autograd aliases plus a side
    channel, so that on the native path the two branch gradients are summed by the dgrad store instead of a separate
    kernel.  Returns ``(x1, x2, fork)``; off the native path it is the identity and ``fork`` is None."""
    if x.is_cuda and not _force_torch() and torch.is_grad_enabled() and x.requires_grad:
        from . import native
        return native.fork(x)
    return x, x, None


def fork_n(x, n: int):
    """n-way version of :func:`fork` for blocks whose input feeds several branches (inception).  Returns
    ``(aliases, fork)``; the branch that is created last passes ``grad_fork=(fork, "deposit_dx")`` (or, for a pooling
    op, ``grad_fork=fork``), the others ``(fork, "accumulate")``."""
    if x.is_cuda and not _force_torch() and torch.is_grad_enabled() and x.requires_grad:
        from . import native
        return native.fork_n(x, n)
    return [x] * n, None


def conv_bn_act(x, w, gamma, beta, running_mean, running_var, *, stride=1, padding=0, groups=1,
                training=True, momentum=0.1, eps=1e-5, act="relu", residual=None, sync_group=False,
                grad_fork=None):
    """``grad_fork=(fork, role)`` with role 'accumulate' (this conv's input gradient is added into the gradient the
    other branch deposited), 'deposit_dx' or 'deposit_res' (this op's input / residual gradient is the deposit)."""
    if grad_fork is not None and grad_fork[0] is None:
        grad_fork = None
    if isinstance(padding, (tuple, list)) and len(padding) == 4:
        # (top, left, bottom, right): TF-"same" layers.  The native depthwise kernel takes it as is; everything else pads.
        pt, pl, pb, pr = padding
        dw = groups != 1 and groups == x.shape[1] == w.shape[0] and groups % 8 == 0 and x.is_cuda and not _force_torch()
        if not dw:
            if pt or pl or pb or pr:
                x = torch.nn.functional.pad(x, [pl, pr, pt, pb])
                if x.dim() == 4:
                    x = x.contiguous(memory_format=torch.channels_last)
            padding = 0
    if use_native(x) and groups == 1:
        from . import native
        return native.conv_bn_act(x, w, gamma, beta, running_mean, running_var, stride=stride,
                                  padding=padding, training=training, momentum=momentum, eps=eps,
                                  act=act, residual=residual, sync_group=sync_group, grad_fork=grad_fork)
    if x.is_cuda and not _force_torch() and groups != 1:
        from . import native
        if groups == x.shape[1] == w.shape[0] and w.shape[1] == 1 and w.shape[2] == w.shape[3] and groups % 8 == 0:
            y = native.depthwise_conv2d(x, w, stride, padding)       # native stencil kernel (csrc/ops/depthwise.cu)
        else:
            _library_fallback(f"grouped convolution (groups={groups}, not depthwise)")
            y = ref.conv2d(x, w, None, stride, padding, groups)
        return native.bn_act(y, gamma, beta, running_mean, running_var, training=training,
                             momentum=momentum, eps=eps, act=act, residual=residual,
                             sync_group=sync_group)
    return ref.conv_bn_act(x, w, gamma, beta, running_mean, running_var, stride=stride,
                           padding=padding, groups=groups, training=training, momentum=momentum,
                           eps=eps, act=act, residual=residual, sync_group=sync_group)


def conv_bn_relu_maxpool(x, w, gamma, beta, running_mean, running_var, *, stride=1, padding=0, training=True, momentum=0.1,
                         eps=1e-5, sync_group=False):
    """Classifier stem ``max_pool2d(relu(BN(conv(x))), 3, 2, 1)`` (torchvision resnet.py `conv1 -> bn1 -> relu -> maxpool`).
    Native path: the normalised activation is never materialised (csrc/ops/stem.cu)."""
    if use_native(x):
        from . import native
        return native.conv_bn_relu_maxpool(x, w, gamma, beta, running_mean, running_var, stride=stride, padding=padding,
                                           training=training, momentum=momentum, eps=eps, sync_group=sync_group)
    z = conv_bn_act(x, w, gamma, beta, running_mean, running_var, stride=stride, padding=padding, training=training,
                    momentum=momentum, eps=eps, act="relu", sync_group=sync_group)
    return max_pool2d(z, 3, 2, 1)


def conv2d(x, w, bias=None, stride=1, padding=0, groups=1):
    if use_native(x) and groups == 1:
        from . import native
        return native.conv2d(x, w, bias, stride, padding)
    if x.is_cuda and not _force_torch():
        _library_fallback(f"conv2d with groups={groups}")
    return ref.conv2d(x, w, bias, stride, padding, groups)


def batch_norm(x, gamma, beta, running_mean, running_var, training, momentum, eps, sync_group=False,
               act=None, residual=None):
    if use_native(x) and x.dim() == 4:
        from . import native
        return native.bn_act(x, gamma, beta, running_mean, running_var, training=training,
                             momentum=momentum, eps=eps, act=act or "none", residual=residual,
                             sync_group=sync_group)
    z = ref.batch_norm(x, gamma, beta, running_mean, running_var, training, momentum, eps, sync_group)
    if residual is not None:
        z = z + residual
    return ref.activation(z, act)


def linear(x, w, b=None, act: Optional[str] = None):
    if use_native(x):
        from . import native
        return native.linear(x, w, b, act)
    return ref.linear(x, w, b, act)


def add_layer_norm(x, delta, gamma, beta, eps=1e-6):
    """Transformer residual step: returns ``(x + delta, LayerNorm(x + delta))``.  On the native path one kernel does both,
    and one kernel the whole backward including the gradient sum over the residual stream."""
    if x.is_cuda and not _force_torch() and x.shape[-1] % 8 == 0 and x.shape[-1] <= 1024:
        require_lib()
        from . import native
        return native.add_layer_norm(x, delta, gamma, beta, eps)
    s = x + delta
    return s, layer_norm(s, gamma, beta, eps)


def activation(x, act):
    if act in (None, "none"):
        return x
    if use_native(x):
        from . import native
        return native.activation(x, act)
    return ref.activation(x, act)


def max_pool2d(x, kernel_size, stride, padding=0, ceil_mode=False, grad_fork=None):
    if use_native(x) and not ceil_mode:
        from . import native
        return native.pool2d(x, kernel_size, stride, padding, mode="max", grad_fork=grad_fork)
    if x.is_cuda and not _force_torch():
        _library_fallback("max_pool2d(ceil_mode=True)")
    return ref.max_pool2d(x, kernel_size, stride, padding, ceil_mode)


def avg_pool2d(x, kernel_size, stride, padding=0, count_include_pad=True, grad_fork=None):
    if use_native(x) and count_include_pad:
        from . import native
        return native.pool2d(x, kernel_size, stride, padding, mode="avg", grad_fork=grad_fork)
    if x.is_cuda and not _force_torch():
        _library_fallback("avg_pool2d(count_include_pad=False)")
    return ref.avg_pool2d(x, kernel_size, stride, padding, count_include_pad)


def channel_scale(x, s, sample_mask=None):
    """``x * s[:, :, None, None]`` (squeeze-excite gate, s = [N, C]) times an optional per-sample factor ``sample_mask`` [N]
    (stochastic depth: mask / keep_prob).  Native path: one kernel forward, one backward (dx and the [N, C] gate gradient)."""
    if use_native(x) and x.dim() == 4 and x.shape[1] % 8 == 0:
        from . import native
        return native.channel_scale(x, s, sample_mask)
    out = x * s.to(x.dtype)[:, :, None, None]
    if sample_mask is not None:
        out = out * sample_mask.to(x.dtype).view(-1, 1, 1, 1)
    return out


def global_avg_pool(x):
    if use_native(x):
        from . import native
        return native.global_avg_pool(x)
    return ref.global_avg_pool(x)


def cross_entropy(logits, labels, weight=None):
    if use_native(logits):
        from . import native
        return native.cross_entropy(logits, labels, weight)
    return ref.cross_entropy(logits, labels, weight)


def layer_norm(x, gamma, beta, eps=1e-6):
    if use_native(x):
        from . import native
        return native.layer_norm(x, gamma, beta, eps)
    return ref.layer_norm(x, gamma, beta, eps)


def attention(q, k, v, scale=None):
    if use_native(q):
        from . import native
        _library_fallback(f"attention with head dim {q.shape[-1]} / {q.shape[-2]} tokens (native: 64 / <= 256)")
        return native.attention(q, k, v, scale)
    return ref.attention(q, k, v, scale)


def attention_packed(qkv, batch: int, seq: int, heads: int):
    """Self-attention on the packed QKV projection ``[batch*seq, 3*heads*dh]`` -> ``[batch*seq, heads*dh]``.
    CUDA + head dim 64 + seq <= 256 runs the tcgen05 kernels; everything else the generic path."""
    d3 = qkv.shape[-1]
    dh = d3 // (3 * heads)
    if use_native(qkv) and dh == 64:
        from . import native
        if native.attention_packed_supported(qkv, seq, heads):
            return native.attention_packed(qkv, batch, seq, heads)
    q = qkv.view(batch, seq, 3, heads, dh).permute(2, 0, 3, 1, 4)
    o = attention(q[0], q[1], q[2])
    return o.permute(0, 2, 1, 3).reshape(batch * seq, heads * dh)


def dropout(x, p, training):
    if not training or p == 0.0:
        return x
    if use_native(x):
        from . import native
        return native.dropout(x, p)
    return ref.dropout(x, p, training)


def synthetic_batch(batch, channels, height, width, num_classes, seed, step, device="cpu",
                    dtype=torch.float32):
    dev = torch.device(device)
    if dev.type == "cuda" and not _force_torch():
        require_lib()
        from . import native
        return native.synthetic_batch(batch, channels, height, width, num_classes, seed, step, dev, dtype)
    return ref.synthetic_batch(batch, channels, height, width, num_classes, seed, step, dev, dtype)


def normalize_u8(images_u8, dtype=torch.float32):
    if use_native(images_u8):
        from . import native
        return native.normalize_u8(images_u8, dtype)
    return ref.normalize_u8(images_u8, dtype)


def concat_channels(xs):
    """torch.cat(xs, dim=1) for NHWC activations (inception blocks)."""
    xs = list(xs)
    if use_native(xs[0]) and all(x.dtype == torch.bfloat16 for x in xs):
        from . import native
        return native.concat_channels(xs)
    out = torch.cat(xs, 1)
    return out.contiguous(memory_format=torch.channels_last) if out.dim() == 4 else out


def top1_hits(logits, labels):
    """Per-sample top-1 hit as int64 [N] (reference utils.py:25-27)."""
    if logits.dim() == 2 and use_native(logits):
        from . import native
        return native.top1_hits(logits, labels)
    return 1 * (logits.argmax(dim=1) == labels)
