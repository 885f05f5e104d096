#!/usr/bin/env python
"""Run one convolution shape a few times (target for `ncu --set full`).
    python benchmarks/one_conv.py --shape 256,56,56,64,256,1,1,1,0 --op fprop --iters 3"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.ops import native  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shape", default="256,56,56,64,256,1,1,1,0")
ap.add_argument("--op", default="fprop", choices=["fprop", "dgrad", "wgrad"])
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--no-stats", action="store_true")
a = ap.parse_args()
N, H, W, Cin, Cout, R, S, stride, pad = [int(v) for v in a.shape.split(",")]
dev = torch.device("cuda")
OH, OW = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
x = torch.randn(N, Cin, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
w = torch.randn(Cout, Cin, R, S, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
y = torch.empty(N, Cout, OH, OW, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
dy = torch.randn_like(y)
dx = torch.empty_like(x)
dw = torch.zeros(Cout, Cin, R, S, device=dev).contiguous(memory_format=torch.channels_last)
stats = None if a.no_stats else native.stat_buffer(Cout, dev)
geo = (N, H, W, Cin, Cout, R, S, stride, pad, pad, OH, OW)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.iters + 1)]
ev[0].record()
for i in range(a.iters):
    if a.op == "fprop":
        native.call("picd_conv_fprop", x, w, y, stats, None, 0, *geo)
    elif a.op == "dgrad":
        native.call("picd_conv_dgrad", dy, w, dx, *geo)
    else:
        native.call("picd_conv_wgrad", x, dy, dw, *geo)
    ev[i + 1].record()
torch.cuda.synchronize()
print(a.op, a.shape, "ms per launch:", [round(ev[i].elapsed_time(ev[i + 1]), 4) for i in range(a.iters)])
