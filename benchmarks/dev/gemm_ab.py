"""A/B of the GEMM launch modes (PICD_QUAD / PICD_FREE_SCHED are read once per process: run once per setting)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from picd_b200.ops import native
dev = torch.device("cuda")
def t(fn, iters=20, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
for (M, N, K) in [(8192, 8192, 8192), (50432, 3072, 768), (50432, 768, 3072), (50432, 2304, 768), (25216, 3072, 768), (25216, 768, 3072)]:
    a = torch.randn(M, K, device=dev).to(torch.bfloat16)
    b = torch.randn(N, K, device=dev).to(torch.bfloat16)
    d = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    ms = t(lambda: native.call("picd_gemm_nt", a, b, d, None, 0, M, N, K))
    print(json.dumps({"mode": os.environ.get("PICD_QUAD", "default"), "shape": [M, N, K], "ms": round(ms, 4), "tflops": round(2.0 * M * N * K / ms / 1e9, 1)}), flush=True)
