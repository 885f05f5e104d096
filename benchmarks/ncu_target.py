#!/usr/bin/env python
"""One launch of every hot kernel on a representative shape (two passes in a plain run; one with PICD_NCU_SINGLE_PASS=1,
which is what `ncu --set full -k regex:"conv_|bn_|attn_|layernorm" -c 40` captures after the plain run has exited 0:
B200_PROFILING.md -- profile only what has run clean without ncu)."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.ops import native, require_lib  # noqa: E402

dev = torch.device("cuda")
bf, cl = torch.bfloat16, torch.channels_last
lib = require_lib()
lib.picd_launch_count.restype = ctypes.c_ulonglong


def conv_case(N, H, Cin, Cout, R, stride):
    pad = R // 2
    OH = (H + 2 * pad - R) // stride + 1
    x = torch.randn(N, Cin, H, H, device=dev).to(bf).contiguous(memory_format=cl)
    w = torch.randn(Cout, Cin, R, R, device=dev).to(bf).contiguous(memory_format=cl)
    y = torch.empty(N, Cout, OH, OH, device=dev, dtype=bf).contiguous(memory_format=cl)
    dy = torch.randn_like(y)
    dw = torch.zeros(Cout, Cin, R, R, device=dev).contiguous(memory_format=cl)
    dx = torch.empty_like(x)
    return dict(x=x, w=w, y=y, dy=dy, dw=dw, dx=dx, geo=(N, H, H, Cin, Cout, R, R, stride, pad, pad, OH, OH),
                stats=torch.zeros(16 * 2 * Cout, device=dev), yup=torch.randn_like(x), coef=torch.rand(4, Cin, device=dev) + 0.5,
                sums=torch.zeros(16 * 2 * Cin, device=dev))


c33 = conv_case(256, 56, 64, 64, 3, 1)          # halo mode, resident weights
c11 = conv_case(256, 56, 64, 256, 1, 1)         # short-K 1x1
c28 = conv_case(256, 28, 128, 128, 3, 1)
c14 = conv_case(256, 14, 256, 256, 3, 1)        # CTA pairs
P, C = 256 * 56 * 56, 256
yv = torch.randn(P, C, device=dev).to(bf)
dz, out = torch.randn_like(yv), torch.empty_like(yv)
coef = torch.rand(4, C, device=dev) + 0.5
sums = torch.zeros(16 * 2 * C, device=dev)
dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
B, S, H = 128, 197, 12
qkv = torch.randn(B * S, 3 * H * 64, device=dev).to(bf)
o = torch.empty(B * S, H * 64, device=dev, dtype=bf)
do = torch.randn_like(o)
lse, delta = torch.empty(B, H, S, device=dev), torch.empty(B, H, S, device=dev)
dqkv = torch.empty_like(qkv)
M, Nn, K = 25216, 3072, 768
gx, gw = torch.randn(M, K, device=dev).to(bf), torch.randn(Nn, K, device=dev).to(bf)
gy = torch.empty(M, Nn, device=dev, dtype=bf)
gdw, gdb = torch.zeros(Nn, K, device=dev), torch.zeros(Nn, device=dev)
lx, ld = torch.randn(M, K, device=dev).to(bf), torch.randn(M, K, device=dev).to(bf)
lo = torch.empty_like(lx)
lstat = torch.rand(2, M, device=dev)
lgm, lg, lb = torch.rand(K, device=dev), torch.zeros(K, device=dev), torch.zeros(K, device=dev)


def one_pass():
    native.call("picd_conv_fprop", c33["x"], c33["w"], c33["y"], c33["stats"], None, 0, *c33["geo"])
    native.call("picd_conv_fprop", c11["x"], c11["w"], c11["y"], c11["stats"], None, 0, *c11["geo"])
    native.call("picd_conv_fprop", c14["x"], c14["w"], c14["y"], c14["stats"], None, 0, *c14["geo"])
    native.call("picd_conv_dgrad", c28["dy"], c28["w"], c28["dx"], *c28["geo"])
    native.call("picd_conv_dgrad_bnreduce", c11["dy"], c11["w"], c11["dx"], *c11["geo"], c11["yup"], c11["coef"][0], c11["coef"][1],
                c11["coef"][2], c11["coef"][3], 1, c11["sums"])
    native.call("picd_conv_wgrad", c28["x"], c28["dy"], c28["dw"], *c28["geo"])
    native.call("picd_bn_apply", yv, coef[0], coef[1], None, 1, out, P, C)
    native.call("picd_bn_bwd_reduce", dz, yv, None, coef[0], coef[1], coef[2], coef[3], 1, sums, P, C)
    native.call("picd_bn_bwd_apply", dz, yv, None, coef[0], coef[1], coef[2], coef[3], 1, sums, 16, 1.0 / P, sums, 16, dg, db, out,
                None, P, C)
    native.call("picd_attention_fwd", qkv, o, lse, B, S, H)
    native.call("picd_attention_bwd", qkv, o, do, lse, delta, dqkv, B, S, H)
    native.call("picd_gemm_nt", gx, gw, gy, None, 0, M, Nn, K)
    native.call("picd_gemm_tn_f32_bias", gy, gx, gdw, gdb, M, Nn, K)
    native.call("picd_layernorm_bwd_add", lx, ld, lgm, lstat[0], lstat[1], ld, lo, lg, lb, M, K)


n0 = lib.picd_launch_count()
one_pass()
torch.cuda.synchronize()
n1 = lib.picd_launch_count()
if os.environ.get("PICD_NCU_SINGLE_PASS", "0") != "1":       # under ncu one pass is enough (every kernel is replayed anyway)
    one_pass()
    torch.cuda.synchronize()
print(f"launches_per_pass={n1 - n0}")
