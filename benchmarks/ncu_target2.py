#!/usr/bin/env python
"""One launch of every kernel added late in round 2 (fused stem, LayerNorm with shared-memory partials, depthwise strips,
squeeze-excite gate, the blocked GEMM tile walk) on its production shape; `ncu --set full -k regex:"maxpool|layernorm|dw_|
channel_scale|conv_gemm" -c 20` captures it after the plain run has exited 0."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.ops import native  # noqa: E402

dev = torch.device("cuda")
bf, cl = torch.bfloat16, torch.channels_last
# ResNet stem at batch 256
N, C, H, OH = 256, 64, 112, 56
y = torch.randn(N, C, H, H, device=dev).to(bf).contiguous(memory_format=cl)
coef = torch.rand(4, C, device=dev) + 0.5
pooled = torch.empty(N, C, OH, OH, device=dev, dtype=bf).contiguous(memory_format=cl)
arg = torch.empty(N, OH, OH, C, dtype=torch.uint8, device=dev)
dpool = torch.randn_like(pooled)
sums = torch.zeros(16 * 2 * C, device=dev)
dy = torch.empty_like(y)
dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
# LayerNorm at the ViT-B/16 shape
M, D = 25216, 768
lx, ld = torch.randn(M, D, device=dev).to(bf), torch.randn(M, D, device=dev).to(bf)
lsum, lyo, ldx = torch.empty_like(lx), torch.empty_like(lx), torch.empty_like(lx)
lmean, lrstd = torch.empty(M, device=dev), torch.empty(M, device=dev)
lg, lb = torch.ones(D, device=dev), torch.zeros(D, device=dev)
ldg, ldb = torch.zeros(D, device=dev), torch.zeros(D, device=dev)
# depthwise 5x5 stride 1 (EfficientNet-B3 stage 3) and the squeeze-excite gate on the same tensor
dN, dC, dH, K = 64, 288, 38, 5
dxin = torch.randn(dN, dC, dH, dH, device=dev).to(bf).contiguous(memory_format=cl)
dwt = torch.randn(K * K, dC, device=dev)
dyo, ddx = torch.empty_like(dxin), torch.empty_like(dxin)
ddw = torch.zeros(K * K, dC, device=dev)
gate = torch.rand(dN, dC, device=dev).to(bf)
gds = torch.zeros(dN, dC, device=dev)
# big GEMM
G = 8192
ga, gb = torch.randn(G, G, device=dev).to(bf), torch.randn(G, G, device=dev).to(bf)
gd = torch.empty(G, G, device=dev, dtype=bf)


def one_pass():
    native.call("picd_bn_relu_maxpool_fwd", y, coef[0], coef[1], pooled, arg, N, H, H, C, OH, OH)
    native.call("picd_bn_relu_maxpool_bwd_reduce", dpool, arg, y, coef[0], coef[1], coef[2], coef[3], sums, N, H, H, C, OH, OH)
    native.call("picd_bn_relu_maxpool_bwd_apply", dpool, arg, y, coef[0], coef[1], coef[2], coef[3], sums, 16, 1.0 / (N * H * H), None,
                sums, 16, dg, db, dy, N, H, H, C, OH, OH)
    native.call("picd_add_layernorm_fwd", lx, ld, lg, lb, lsum, lyo, lmean, lrstd, M, D, 1e-6)
    native.call("picd_layernorm_bwd_add", ld, lx, lg, lmean, lrstd, lsum, ldx, ldg, ldb, M, D)
    native.call("picd_depthwise_fprop", dxin, dwt, dyo, dN, dH, dH, dC, dH, dH, K, 1, 2, 2)
    native.call("picd_depthwise_dgrad", dyo, dwt, ddx, dN, dH, dH, dC, dH, dH, K, 1, 2, 2)
    native.call("picd_depthwise_wgrad", dxin, dyo, ddw, dN, dH, dH, dC, dH, dH, K, 1, 2, 2)
    native.call("picd_channel_scale_fwd", dxin, gate, None, dyo, dN, dH * dH, dC)
    native.call("picd_channel_scale_bwd", dyo, dxin, gate, None, ddx, gds, dN, dH * dH, dC)
    native.call("picd_gemm_nt", ga, gb, gd, None, 0, G, G, G)


one_pass()
torch.cuda.synchronize()
if os.environ.get("PICD_NCU_SINGLE_PASS", "0") != "1":
    one_pass()
    torch.cuda.synchronize()
print("ok")
