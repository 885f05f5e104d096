#!/usr/bin/env python
"""dgrad + BatchNorm-backward reduction, separate kernels vs the fused epilogue (picd_conv_dgrad_bnreduce), for every
ResNet-50 layer whose output gradient is produced by exactly one dgrad (bn1 / bn2 of each bottleneck).  CUDA events,
inputs larger than L2 at batch 256.  Writes gpurun_out/fused_reduce_layers_b<batch>.txt."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.ops import native  # noqa: E402

dev = torch.device("cuda")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
# (H of dx, Cin (= channels of dx / upstream BN), Cout, R, stride, count)
layers = [(56, 64, 64, 3, 1, 3), (56, 64, 256, 1, 1, 3), (56, 128, 128, 3, 2, 1), (28, 128, 128, 3, 1, 3), (28, 128, 512, 1, 1, 4),
          (28, 256, 256, 3, 2, 1), (14, 256, 256, 3, 1, 5), (14, 256, 1024, 1, 1, 6), (14, 512, 512, 3, 2, 1),
          (7, 512, 512, 3, 1, 2), (7, 512, 2048, 1, 1, 3)]


def timeit(fn, iters=6):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


lines, tot = [], {}
for H, Cin, Cout, R, stride, cnt in layers:
    pad = R // 2
    OH = (H + 2 * pad - R) // stride + 1
    dy = torch.randn(N, Cout, OH, OH, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = torch.randn(Cout, Cin, R, R, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dx = torch.empty(N, Cin, H, H, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = torch.randn_like(dx)
    coef = torch.rand(4, Cin, device=dev) + 0.5
    sums = torch.zeros(16 * 2 * Cin, device=dev)
    geo = (N, H, H, Cin, Cout, R, R, stride, pad, pad, OH, OH)
    P = N * H * H
    t_d = timeit(lambda: native.call("picd_conv_dgrad", dy, w, dx, *geo))
    t_r = timeit(lambda: native.call("picd_bn_bwd_reduce", dx, y, None, coef[0], coef[1], coef[2], coef[3], 1, sums, P, Cin))
    res = {"dgrad": t_d, "reduce": t_r}
    for yb in (1, 2, 3):
        os.environ["PICD_YBUFS"] = str(yb)
        try:
            res[f"fused_y{yb}"] = timeit(lambda: native.call("picd_conv_dgrad_bnreduce", dy, w, dx, *geo, y, coef[0], coef[1], coef[2],
                                                             coef[3], 1, sums))
        except Exception as e:  # noqa: BLE001
            res[f"fused_y{yb}"] = float("nan")
    os.environ.pop("PICD_YBUFS")
    res["fused_auto"] = timeit(lambda: native.call("picd_conv_dgrad_bnreduce", dy, w, dx, *geo, y, coef[0], coef[1], coef[2], coef[3],
                                                   1, sums))
    gb = dx.numel() * 2 / 1e9
    lines.append(f"dx {H:3d}x{H:<3d} C={Cin:<4d} <- {Cout:<5d} {R}x{R} s{stride} x{cnt} ({gb:5.3f} GB) | dgrad {t_d:6.3f} + reduce {t_r:6.3f} = "
                 f"{t_d + t_r:6.3f} ms | fused y1 {res['fused_y1']:6.3f} y2 {res['fused_y2']:6.3f} y3 {res['fused_y3']:6.3f} auto {res['fused_auto']:6.3f} ms")
    for k, v in res.items():
        tot[k] = tot.get(k, 0.0) + v * cnt
    del dy, w, dx, y
    torch.cuda.empty_cache()
lines.append("TOTAL per step: " + "  ".join(f"{k} {v:.3f}" for k, v in tot.items()) + " ms")
print("\n".join(lines))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", f"fused_reduce_layers_b{N}.txt"), "w").write("\n".join(lines) + "\n")
