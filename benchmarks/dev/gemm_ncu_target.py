"""One big GEMM a few times (ncu target): PICD_QUAD selects the launch mode."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from picd_b200.ops import native
dev = torch.device("cuda")
M = N = K = 8192
a = torch.randn(M, K, device=dev).to(torch.bfloat16)
b = torch.randn(N, K, device=dev).to(torch.bfloat16)
d = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
for _ in range(3):
    native.call("picd_gemm_nt", a, b, d, None, 0, M, N, K)
for _ in range(3):
    torch.matmul(a, b.t(), out=d)
torch.cuda.synchronize()
print("ok")
