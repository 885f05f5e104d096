#!/usr/bin/env python
"""Per-shape timings of the BatchNorm kernels (achieved GB/s against the measured HBM copy bandwidth)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.ops import native  # noqa: E402

dev = torch.device("cuda")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
shapes = [(112, 64, 1), (56, 64, 6), (56, 256, 4), (56, 128, 1), (28, 128, 7), (28, 512, 5), (28, 256, 1), (14, 256, 11),
          (14, 1024, 7), (14, 512, 1), (7, 512, 5), (7, 2048, 4)]


def timeit(fn, iters=5):
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


tot = [0, 0, 0, 0]
lines = []
for H, C, cnt in shapes:
    P = N * H * H
    y = torch.randn(P, C, device=dev).to(torch.bfloat16)
    dz = torch.randn_like(y)
    z = torch.randn_like(y)
    out = torch.empty_like(y)
    dres = torch.empty_like(y)
    coef = torch.rand(4, C, device=dev) + 0.5
    sums = native.stat_buffer(C, dev)
    dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    gb = y.numel() * 2 / 1e9
    t_st = timeit(lambda: native.call("picd_bn_stats", y, sums, P, C))
    t_ap = timeit(lambda: native.call("picd_bn_apply", y, coef[0], coef[1], None, 1, out, P, C))
    t_rd = timeit(lambda: native.call("picd_bn_bwd_reduce", dz, y, None, coef[0], coef[1], coef[2], coef[3], 1, sums, P, C))
    t_ba = timeit(lambda: native.call("picd_bn_bwd_apply", dz, y, None, coef[0], coef[1], coef[2], coef[3], 1, sums, 16, 1.0 / P,
                                      sums, 16, dg, db, out, None, P, C))
    lines.append(f"{H:3d}x{H:<3d} C={C:<5d} x{cnt:<2d} {gb:6.3f} GB | stats {t_st:6.3f} ms {gb / t_st * 1e3:6.0f} GB/s | apply {t_ap:6.3f} ms "
                 f"{2 * gb / t_ap * 1e3:6.0f} GB/s | bwd_reduce {t_rd:6.3f} ms {2 * gb / t_rd * 1e3:6.0f} GB/s | bwd_apply {t_ba:6.3f} ms "
                 f"{3 * gb / t_ba * 1e3:6.0f} GB/s")
    for i, t in enumerate((t_st, t_ap, t_rd, t_ba)):
        tot[i] += t * cnt
    del y, dz, z, out, dres
    torch.cuda.empty_cache()
lines.append(f"TOTAL weighted: stats {tot[0]:.2f} apply {tot[1]:.2f} bwd_reduce {tot[2]:.2f} bwd_apply {tot[3]:.2f} ms")
print("\n".join(lines))
open(os.path.join(ROOT, "gpurun_out", f"bn_layers_b{N}.txt"), "w").write("\n".join(lines) + "\n")
