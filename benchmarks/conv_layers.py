#!/usr/bin/env python
"""Per-layer convolution timings: our tcgen05 implicit-GEMM kernels vs the library (cuDNN through torch, NHWC bf16)
for every distinct conv shape of a model at a given batch.  CUDA events, 3 warm-ups, inputs >> L2 at batch 256.
Writes gpurun_out/conv_layers_<model>_b<batch>.txt."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from picd_b200.models import Classifier  # noqa: E402
from picd_b200.models.layers import Conv2d  # noqa: E402
from picd_b200.ops import native  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--size", type=int, default=224)
args = ap.parse_args()
dev = torch.device("cuda")

# collect (shape -> count) by running a tiny forward with hooks on the CPU oracle path
shapes = {}
m = Classifier(args.model, 7, image_size=args.size)
hooks = []
for mod in m.modules():
    if isinstance(mod, Conv2d):
        def pre(mod, inp, _shapes=shapes):
            pass
import picd_b200.ops as ops  # noqa: E402
orig = ops.conv_bn_act
def spy(x, w, *a, stride=1, padding=0, groups=1, **kw):
    key = (x.shape[2], x.shape[3], w.shape[1], w.shape[0], w.shape[2], w.shape[3],
           stride[0] if isinstance(stride, tuple) else stride, tuple(padding) if isinstance(padding, (tuple, list)) else (padding, padding))
    shapes[key] = shapes.get(key, 0) + 1
    return orig(x, w, *a, stride=stride, padding=padding, groups=groups, **kw)
ops.conv_bn_act = spy
import picd_b200.models.layers as L  # noqa: E402
m.eval()
with torch.no_grad():
    m(torch.zeros(1, 3, args.size, args.size))
ops.conv_bn_act = orig


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


lines = []
tot = {"f": 0, "d": 0, "w": 0, "cf": 0, "cd": 0, "cw": 0}
N = args.batch
for (H, W, Cin, Cout, R, S, stride, pad), count in sorted(shapes.items(), key=lambda kv: -kv[0][0]):
    if Cin % 8 != 0:
        continue
    OH, OW = (H + 2 * pad[0] - R) // stride + 1, (W + 2 * pad[1] - S) // stride + 1
    x = torch.randn(N, Cin, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = torch.randn(Cout, Cin, R, S, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = torch.empty(N, Cout, OH, OW, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dy = torch.randn_like(y)
    dx = torch.empty_like(x)
    dw = torch.zeros(Cout, Cin, R, S, device=dev).contiguous(memory_format=torch.channels_last)
    stats = native.stat_buffer(Cout, dev)
    geo = (N, H, W, Cin, Cout, R, S, stride, pad[0], pad[1], OH, OW)
    tf = timeit(lambda: native.call("picd_conv_fprop", x, w, y, stats, None, 0, *geo))
    td = timeit(lambda: native.call("picd_conv_dgrad", dy, w, dx, *geo))
    tw = timeit(lambda: native.call("picd_conv_wgrad", x, dy, dw, *geo))
    cf = timeit(lambda: F.conv2d(x, w, None, stride, pad))
    cd = timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [stride, stride], list(pad), [1, 1], False, [0, 0], 1, [True, False, False]))
    cw = timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [stride, stride], list(pad), [1, 1], False, [0, 0], 1, [False, True, False]))
    flops = 2.0 * N * OH * OW * Cout * Cin * R * S
    byts = 2.0 * (x.numel() + y.numel())
    lines.append(f"{H:4d}x{W:<4d} {Cin:5d}->{Cout:<5d} {R}x{S} s{stride} x{count:<2d} | fprop {tf:7.3f} (cudnn {cf:7.3f}) "
                 f"dgrad {td:7.3f} ({cd:7.3f}) wgrad {tw:7.3f} ({cw:7.3f}) ms | fprop {flops / tf / 1e9:7.1f} TF/s "
                 f"{byts / tf / 1e6:7.1f} GB/s")
    for k, v in (("f", tf), ("d", td), ("w", tw), ("cf", cf), ("cd", cd), ("cw", cw)):
        tot[k] += v * count
    del x, w, y, dy, dx, dw
    torch.cuda.empty_cache()
lines.append(f"TOTAL per step (weighted by layer count): ours fprop {tot['f']:.2f} dgrad {tot['d']:.2f} wgrad {tot['w']:.2f} ms; "
             f"library fprop {tot['cf']:.2f} dgrad {tot['cd']:.2f} wgrad {tot['cw']:.2f} ms")
path = os.path.join(ROOT, "gpurun_out", f"conv_layers_{args.model}_b{args.batch}.txt")
os.makedirs(os.path.dirname(path), exist_ok=True)
open(path, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
