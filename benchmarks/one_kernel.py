#!/usr/bin/env python
"""Launch one kernel a few times on a representative shape (target for `ncu --set full -k regex:<name>`).
    python benchmarks/one_kernel.py --what wgrad|fprop3x3|bn_bwd_reduce|bn_bwd_apply|attn_fwd|attn_bwd|gemm"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.ops import native  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--what", required=True)
ap.add_argument("--iters", type=int, default=3)
a = ap.parse_args()
dev = torch.device("cuda")
bf = torch.bfloat16
cl = torch.channels_last


def conv_tensors(N, H, W, Cin, Cout, R, S, stride, pad):
    OH, OW = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    x = torch.randn(N, Cin, H, W, device=dev).to(bf).contiguous(memory_format=cl)
    w = torch.randn(Cout, Cin, R, S, device=dev).to(bf).contiguous(memory_format=cl)
    y = torch.empty(N, Cout, OH, OW, device=dev, dtype=bf).contiguous(memory_format=cl)
    dy = torch.randn_like(y)
    dw = torch.zeros(Cout, Cin, R, S, device=dev).contiguous(memory_format=cl)
    return x, w, y, dy, dw, (N, H, W, Cin, Cout, R, S, stride, pad, pad, OH, OW)


if a.what in ("wgrad", "fprop3x3", "dgrad3x3"):
    x, w, y, dy, dw, geo = conv_tensors(256, 28, 28, 128, 128, 3, 3, 1, 1)
    stats = native.stat_buffer(128, dev)
    dx = torch.empty_like(x)
    fn = {"wgrad": lambda: native.call("picd_conv_wgrad", x, dy, dw, *geo),
          "fprop3x3": lambda: native.call("picd_conv_fprop", x, w, y, stats, None, 0, *geo),
          "dgrad3x3": lambda: native.call("picd_conv_dgrad", dy, w, dx, *geo)}[a.what]
elif a.what.startswith("bn_"):
    P, C = 256 * 56 * 56, 256
    yv = torch.randn(P, C, device=dev).to(bf)
    dz = torch.randn_like(yv)
    out = torch.empty_like(yv)
    coef = torch.rand(4, C, device=dev) + 0.5
    sums = native.stat_buffer(C, dev)
    dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    fn = {"bn_bwd_reduce": lambda: native.call("picd_bn_bwd_reduce", dz, yv, None, coef[0], coef[1], coef[2], coef[3], 1, sums, P, C),
          "bn_bwd_apply": lambda: native.call("picd_bn_bwd_apply", dz, yv, None, coef[0], coef[1], coef[2], coef[3], 1, sums, 16,
                                              1.0 / P, sums, 16, dg, db, out, None, P, C),
          "bn_apply": lambda: native.call("picd_bn_apply", yv, coef[0], coef[1], None, 1, out, P, C)}[a.what]
elif a.what.startswith("attn"):
    B, S, H = 128, 197, 12
    qkv = torch.randn(B * S, 3 * H * 64, device=dev).to(bf)
    o = torch.empty(B * S, H * 64, device=dev, dtype=bf)
    do = torch.randn_like(o)
    lse = torch.empty(B, H, S, device=dev)
    delta = torch.empty_like(lse)
    dqkv = torch.empty_like(qkv)
    native.call("picd_attention_fwd", qkv, o, lse, B, S, H)
    fn = {"attn_fwd": lambda: native.call("picd_attention_fwd", qkv, o, lse, B, S, H),
          "attn_bwd": lambda: native.call("picd_attention_bwd", qkv, o, do, lse, delta, dqkv, B, S, H)}[a.what]
elif a.what == "gemm":
    M, N, K = 25216, 3072, 768
    x = torch.randn(M, K, device=dev).to(bf)
    w = torch.randn(N, K, device=dev).to(bf)
    y = torch.empty(M, N, device=dev, dtype=bf)
    bias = torch.randn(N, device=dev)
    fn = lambda: native.call("picd_gemm_nt", x, w, y, bias, 0, M, N, K)
else:
    raise SystemExit("unknown --what")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.iters + 1)]
ev[0].record()
for i in range(a.iters):
    fn()
    ev[i + 1].record()
torch.cuda.synchronize()
print(a.what, "ms per launch:", [round(ev[i].elapsed_time(ev[i + 1]), 4) for i in range(a.iters)])
