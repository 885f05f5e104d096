#!/usr/bin/env python
"""ViT-B/16 GEMM shapes (M = batch*197): our tcgen05 kernels vs cuBLAS, forward (+bias), dgrad, wgrad."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.ops import native  # noqa: E402

dev = torch.device("cuda")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
M = B * 197


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


lines = []
for (N, K) in [(2304, 768), (768, 768), (3072, 768), (768, 3072)]:
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    w = torch.randn(N, K, device=dev).to(torch.bfloat16)
    bias = torch.randn(N, device=dev)
    y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(M, N, device=dev).to(torch.bfloat16)
    dx = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
    dw = torch.zeros(N, K, device=dev)
    fl = 2.0 * M * N * K
    t_f = timeit(lambda: native.call("picd_gemm_nt", x, w, y, bias, 0, M, N, K))
    t_f0 = timeit(lambda: native.call("picd_gemm_nt", x, w, y, None, 0, M, N, K))
    t_d = timeit(lambda: native.call("picd_gemm_nn", dy, w, dx, M, N, K))
    t_w = timeit(lambda: native.call("picd_gemm_tn_f32", dy, x, dw, M, N, K))
    c_f = timeit(lambda: torch.addmm(bias.to(torch.bfloat16), x, w.t()))
    c_d = timeit(lambda: torch.matmul(dy, w))
    c_w = timeit(lambda: torch.matmul(dy.t(), x))
    lines.append(f"M={M} N={N:5d} K={K:5d} | fwd+bias {t_f:.3f} ms {fl / t_f / 1e9:6.0f} TF/s (no bias {t_f0:.3f}; cuBLAS {c_f:.3f}) | "
                 f"dgrad {t_d:.3f} ms {fl / t_d / 1e9:6.0f} TF/s (cuBLAS {c_d:.3f}) | wgrad {t_w:.3f} ms {fl / t_w / 1e9:6.0f} TF/s (cuBLAS {c_w:.3f})")
print("\n".join(lines))
open(os.path.join(ROOT, "gpurun_out", f"gemm_shapes_b{B}.txt"), "w").write("\n".join(lines) + "\n")
