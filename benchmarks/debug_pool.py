import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from picd_b200.ops import native
dev = torch.device("cuda")
torch.manual_seed(0)
for (N, C, H) in [(1, 8, 5), (2, 64, 17)]:
    x = torch.randn(N, C, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xn = x.clone().requires_grad_()
    y = native.pool2d(xn, 3, 1, 1, "avg")
    dy = torch.randn_like(y)
    y.backward(dy)
    xr = x.float().requires_grad_()
    yr = F.avg_pool2d(xr, 3, 1, 1)
    yr.backward(dy.float())
    err = (xn.grad.float() - xr.grad).abs()
    print("shape", tuple(x.shape), "max err", err.max().item(), "dy strides", dy.stride(), "y strides", y.stride())
    print("err per (h,w) max over n,c:\n", err.amax((0, 1)))
    # direct kernel call
    dx = torch.empty_like(x)
    native.call("picd_pool_bwd", dy, None, dx, N, H, H, C, H, H, 3, 3, 1, 1, 1)
    print("direct max err", (dx.float() - xr.grad).abs().max().item())
    # independent oracle: backward of a 3x3/s1/p1 average pool == depthwise 3x3 box filter / 9 over dy
    wbox = torch.full((C, 1, 3, 3), 1 / 9.0, device=dev)
    oracle = F.conv2d(dy.float().contiguous(), wbox, None, 1, 1, 1, C)
    print("native vs box-filter oracle", (dx.float() - oracle).abs().max().item(),
          "torch-bwd vs oracle", (xr.grad - oracle).abs().max().item())
    xr2 = x.float().contiguous().requires_grad_()
    F.avg_pool2d(xr2, 3, 1, 1).backward(dy.float().contiguous())
    print("torch NCHW-contiguous bwd vs oracle", (xr2.grad - oracle).abs().max().item())
