"""LayerNorm (+ residual add) forward / backward at the ViT-B/16 shape; CUDA events, rotating buffers larger than L2."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from picd_b200.ops import native
dev = torch.device("cuda")
M, D = int(os.environ.get("M", 25216)), 768
NB = 6
xs = [torch.randn(M, D, device=dev).to(torch.bfloat16) for _ in range(NB)]
ds = [torch.randn(M, D, device=dev).to(torch.bfloat16) for _ in range(NB)]
g = torch.ones(D, device=dev, requires_grad=True); b = torch.zeros(D, device=dev, requires_grad=True)
def t(fn, iters=30, warm=6):
    for i in range(warm): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
lib = native
mean = torch.empty(M, device=dev); rstd = torch.empty(M, device=dev)
y = torch.empty(M, D, device=dev, dtype=torch.bfloat16); so = torch.empty_like(y)
dx = torch.empty_like(y); dg = torch.zeros(D, device=dev); db = torch.zeros(D, device=dev)
f_ms = t(lambda i: lib.call("picd_add_layernorm_fwd", xs[i % NB], ds[i % NB], g, b, so, y, mean, rstd, M, D, 1e-6))
b_ms = t(lambda i: lib.call("picd_layernorm_bwd_add", ds[i % NB], xs[i % NB], g, mean, rstd, xs[(i + 1) % NB], dx, dg, db, M, D))
gb = M * D * 2 * 4 / 1e9
print(json.dumps({"M": M, "D": D, "add_ln_fwd_ms": round(f_ms, 4), "fwd_GBps": round(gb / f_ms * 1e3), "ln_bwd_add_ms": round(b_ms, 4), "bwd_GBps": round(gb / b_ms * 1e3)}))
