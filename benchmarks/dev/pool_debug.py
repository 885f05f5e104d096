import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch, torch.nn.functional as F
from picd_b200 import ops
dev = torch.device("cuda")
for mode in ("avg", "max"):
    for (N, C, H, W, k, s, p) in [(8, 256, 17, 17, 3, 1, 1), (8, 192, 17, 17, 3, 1, 1), (4, 64, 16, 16, 3, 1, 1), (4, 64, 9, 9, 3, 2, 1), (4, 64, 12, 12, 3, 2, 0), (4, 64, 8, 8, 2, 2, 0), (2, 768, 17, 17, 5, 3, 0)]:
        if mode == "max" and k == 5: continue
        x = torch.randn(N, C, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        xn = x.clone().requires_grad_(); xr = x.float().requires_grad_()
        f = ops.avg_pool2d if mode == "avg" else ops.max_pool2d
        fr = F.avg_pool2d if mode == "avg" else F.max_pool2d
        y = f(xn, k, s, p); yr = fr(xr, k, s, p)
        dy = torch.randn_like(yr)
        y.backward(dy.to(torch.bfloat16)); yr.backward(dy.to(torch.bfloat16).float())
        e = lambda a, b: float((a.float() - b).abs().max() / (b.abs().max() + 1e-6))
        print(mode, (N, C, H, W, k, s, p), "fwd", round(e(y, yr), 5), "bwd", round(e(xn.grad, xr.grad), 5))
