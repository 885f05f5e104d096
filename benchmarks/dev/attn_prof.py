import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import ProfilerActivity, profile
from picd_b200.ops import native
dev = torch.device("cuda")
B, S, H = 256, 197, 12
qkv = (torch.randn(B * S, 3 * H * 64, device=dev) * 0.5).to(torch.bfloat16).requires_grad_()
do = torch.randn(B * S, H * 64, device=dev).to(torch.bfloat16)
def fb():
    qkv.grad = None
    native.attention_packed(qkv, B, S, H).backward(do)
for _ in range(3): fb()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5): fb()
    torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:6]:
    print(f"{e.device_time_total / 5 / 1e3:8.3f} ms  x{e.count // 5}  {e.key[:80]}")
