"""Per-parameter cosine table of the compose cases (native / library bf16 vs the fp32 oracle)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
import torch
import torch.nn as nn
import gpu_selftest as G
from picd_b200.models import inception, resnet

dev = torch.device("cuda")
cos = lambda a, b: float(torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0))
which = sys.argv[1] if len(sys.argv) > 1 else "inception"
if which == "inception":
    build = lambda: nn.Sequential(inception.InceptionA(192, 32), inception.InceptionA(256, 64)); shape = (128, 192, 17, 17)
elif which == "inception1":
    build = lambda: nn.Sequential(inception.InceptionA(192, 32)); shape = (128, 192, 17, 17)
else:
    ds = lambda: nn.Sequential(resnet.Conv2d(64, 128, 1, 2), resnet.BatchNorm2d(128))
    build = lambda: nn.Sequential(resnet.BasicBlock(64, 64), resnet.BasicBlock(64, 64), resnet.BasicBlock(64, 128, 2, ds()), resnet.BasicBlock(128, 128)); shape = (128, 64, 28, 28)
o1, g1, dx1 = G._stage_grads(build, shape, "native", dev)
o0, g0, dx0 = G._stage_grads(build, shape, "torch_bf16", dev)
o2, g2, dx2 = G._stage_grads(build, shape, "torch_fp32", dev)
for k in g2:
    print(f"{k:40s} native {cos(g1[k], g2[k]):.4f}  lib {cos(g0[k], g2[k]):.4f}  |g| {float(g2[k].norm()):.3e} native|g| {float(g1[k].norm()):.3e}")
print("input", cos(dx1, dx2), cos(dx0, dx2))
