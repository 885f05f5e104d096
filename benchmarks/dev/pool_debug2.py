import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch, torch.nn.functional as F
from picd_b200 import ops
dev = torch.device("cuda")
torch.set_printoptions(linewidth=200, precision=3)
N, C, H, W, k, s, p = 1, 8, 5, 5, 3, 1, 1
x = torch.randn(N, C, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
xn = x.clone().requires_grad_(); xr = x.float().requires_grad_()
y = ops.avg_pool2d(xn, k, s, p); yr = F.avg_pool2d(xr, k, s, p)
dy = torch.ones_like(yr)
y.backward(dy.to(torch.bfloat16)); yr.backward(dy)
print("native\n", xn.grad[0, 0].float()); print("torch\n", xr.grad[0, 0])
dy = torch.zeros_like(yr); dy[0, :, 0, 0] = 9.0
xn.grad = None; xr.grad = None
y = ops.avg_pool2d(xn, k, s, p); yr = F.avg_pool2d(xr, k, s, p)
y.backward(dy.to(torch.bfloat16)); yr.backward(dy)
print("native\n", xn.grad[0, 0].float()); print("torch\n", xr.grad[0, 0])
