"""ViT-B/16 (driver extension, BASELINE.json config 4) with torchvision's parameter names.

The reference has no attention model; SURVEY.md 0.2 records that ViT has no ``.fc``
so the classifier swaps ``heads.head`` instead.  Patch embedding is a 16x16/s16
convolution, which in NHWC is exactly a GEMM over non-overlapping patches; QKV,
projection and MLP GEMMs run on the tcgen05 kernel with bias/GELU epilogues.
"""
from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn

from .. import ops
from .layers import Conv2d, Dropout, LayerNorm, Linear, prepare_input


class SelfAttention(nn.Module):
    """Parameters named like nn.MultiheadAttention (in_proj_weight/in_proj_bias/out_proj)."""

    def __init__(self, dim, heads):
        super().__init__()
        self.dim, self.heads = dim, heads
        self.in_proj_weight = nn.Parameter(torch.empty(3 * dim, dim))
        nn.init.xavier_uniform_(self.in_proj_weight)
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * dim))
        self.out_proj = Linear(dim, dim)
        nn.init.zeros_(self.out_proj.bias)

    def forward(self, x):
        b, s, d = x.shape
        qkv = ops.linear(x.reshape(b * s, d), self.in_proj_weight, self.in_proj_bias)    # [b*s, 3d], packed q|k|v
        o = ops.attention_packed(qkv, b, s, self.heads)                                    # [b*s, d]
        return self.out_proj(o).view(b, s, d)


class MLPBlock(nn.Sequential):
    def __init__(self, dim, hidden, dropout):
        super().__init__(Linear(dim, hidden, act="gelu"), nn.Identity(), Dropout(dropout),
                         Linear(hidden, dim), Dropout(dropout))
        for i in (0, 3):
            nn.init.xavier_uniform_(self[i].weight)
            nn.init.normal_(self[i].bias, std=1e-6)


class EncoderBlock(nn.Module):
    def __init__(self, dim, heads, hidden, dropout):
        super().__init__()
        self.ln_1 = LayerNorm(dim)
        self.self_attention = SelfAttention(dim, heads)
        self.dropout = Dropout(dropout)
        self.ln_2 = LayerNorm(dim)
        self.mlp = MLPBlock(dim, hidden, dropout)

    def forward(self, x):
        b, s, d = x.shape
        x = x + self.dropout(self.self_attention(self.ln_1(x)))
        y = self.mlp(self.ln_2(x).reshape(b * s, d)).view(b, s, d)
        return x + y

    def forward_fused(self, x, h, next_ln):
        """Same block with the two residual adds fused into the LayerNorms that follow them (``ops.add_layer_norm``):
        ``x`` is the residual stream, ``h = ln_1(x)`` was produced by the previous block's second add, and this block's
        second add is fused with ``next_ln`` (the next block's ``ln_1`` or the encoder's final LayerNorm)."""
        b, s, d = x.shape
        a = self.dropout(self.self_attention(h))
        x, h = ops.add_layer_norm(x, a, self.ln_2.weight, self.ln_2.bias, self.ln_2.eps)
        y = self.mlp(h.reshape(b * s, d)).view(b, s, d)
        return ops.add_layer_norm(x, y, next_ln.weight, next_ln.bias, next_ln.eps)


class Encoder(nn.Module):
    def __init__(self, seq, depth, dim, heads, hidden, dropout):
        super().__init__()
        self.pos_embedding = nn.Parameter(torch.empty(1, seq, dim).normal_(std=0.02))
        self.dropout = Dropout(dropout)
        self.layers = nn.Sequential(OrderedDict(
            (f"encoder_layer_{i}", EncoderBlock(dim, heads, hidden, dropout)) for i in range(depth)))
        self.ln = LayerNorm(dim)

    def forward(self, x):
        x = self.dropout(x + self.pos_embedding.to(x.dtype))
        blocks = list(self.layers)
        h = blocks[0].ln_1(x)
        for i, blk in enumerate(blocks):         # residual adds ride in the LayerNorm kernels (EncoderBlock.forward_fused)
            nxt = blocks[i + 1].ln_1 if i + 1 < len(blocks) else self.ln
            x, h = blk.forward_fused(x, h, nxt)
        return h                                 # = self.ln(residual stream)


class VisionTransformer(nn.Module):
    def __init__(self, image_size=224, patch=16, depth=12, heads=12, dim=768, hidden=3072,
                 num_classes=1000, dropout=0.0):
        super().__init__()
        self.image_size, self.patch, self.hidden_dim = image_size, patch, dim
        self.conv_proj = Conv2d(3, dim, patch, patch, 0, bias=True)
        fan_in = 3 * patch * patch
        nn.init.trunc_normal_(self.conv_proj.weight, std=(1.0 / fan_in) ** 0.5)
        self.class_token = nn.Parameter(torch.zeros(1, 1, dim))
        seq = (image_size // patch) ** 2 + 1
        self.encoder = Encoder(seq, depth, dim, heads, hidden, dropout)
        self.heads = nn.Sequential(OrderedDict(head=Linear(dim, num_classes)))
        nn.init.zeros_(self.heads.head.weight)
        nn.init.zeros_(self.heads.head.bias)

    def forward(self, x):
        x = prepare_input(x)
        n = x.shape[0]
        x = self.conv_proj(x)                                  # [n, dim, h/p, w/p] channels_last
        x = x.permute(0, 2, 3, 1).reshape(n, -1, self.hidden_dim)   # free in NHWC
        x = torch.cat([self.class_token.to(x.dtype).expand(n, -1, -1), x], dim=1)
        x = self.encoder(x)
        return self.heads(x[:, 0])


def vit_b_16(num_classes=1000, image_size=224, **kw):
    return VisionTransformer(image_size, 16, 12, 12, 768, 3072, num_classes, **kw)


def vit_tiny_test(num_classes=10, image_size=32, **kw):
    """A 2-block, 64-wide ViT used by CPU tests."""
    return VisionTransformer(image_size, 8, 2, 2, 64, 128, num_classes, **kw)
