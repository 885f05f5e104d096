import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
import torch, torch.nn as nn
import gpu_selftest as G
from picd_b200.models import inception
from picd_b200 import ops
from picd_b200.ops import native
dev = torch.device("cuda")
cos = lambda a, b: float(torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0))
for cin, pf, hw, n in [(192, 32, 17, 128), (256, 64, 17, 128), (256, 64, 16, 128), (256, 64, 17, 8)]:
    build = lambda: nn.Sequential(inception.InceptionA(cin, pf)); shape = (n, cin, hw, hw)
    o2, g2, dx2 = G._stage_grads(build, shape, "torch_fp32", dev)
    o0, g0, dx0 = G._stage_grads(build, shape, "torch_bf16", dev)
    o1, g1, dx1 = G._stage_grads(build, shape, "native", dev)
    keep = native.fork_n
    native.fork_n = lambda x, k: ([x] * k, None)
    o3, g3, dx3 = G._stage_grads(build, shape, "native", dev)
    native.fork_n = keep
    print(cin, pf, hw, n, "input-grad cos: native(fork)", round(cos(dx1, dx2), 4), "native(nofork)", round(cos(dx3, dx2), 4), "lib", round(cos(dx0, dx2), 4),
          "min param cos native", round(min(cos(g1[k], g2[k]) for k in g2), 4), "lib", round(min(cos(g0[k], g2[k]) for k in g2), 4),
          "|dx| native", float(dx1.norm()), "nofork", float(dx3.norm()), "fp32", float(dx2.norm()))
    d = (dx1 - dx3).abs()
    print("   fork-vs-nofork max abs diff", float(d.max()), "mean", float(d.mean()), "per-channel-block mean diff", [round(float(d[:, c:c + 64].mean()) * 1e6, 3) for c in range(0, cin, 64)])
