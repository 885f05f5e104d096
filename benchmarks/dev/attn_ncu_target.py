import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from picd_b200.ops import native
dev = torch.device("cuda")
B, S, H = 256, 197, 12
qkv = torch.randn(B * S, 3 * H * 64, device=dev).to(torch.bfloat16)
o = torch.empty(B * S, H * 64, device=dev, dtype=torch.bfloat16)
lse = torch.empty(B, H, S, device=dev)
for _ in range(3):
    native.call("picd_attention_fwd", qkv, o, lse, B, S, H)
torch.cuda.synchronize()
print("ok")
