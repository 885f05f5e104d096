import sys, torch
sys.path.insert(0, "/root/repo")
from picd_b200.ops import native
dev = torch.device("cuda")
def timeit(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters
for H, C in [(56, 256), (28, 512), (14, 1024)]:
    P = 256 * H * H
    y = torch.randn(P, C, device=dev).to(torch.bfloat16); dz = torch.randn_like(y); z = torch.randn_like(y)
    out = torch.empty_like(y); dres = torch.empty_like(y)
    coef = torch.rand(4, C, device=dev) + 0.5
    sums = torch.zeros(16 * 2 * C, device=dev)
    mask = torch.randint(0, 256, (P * C // 8 + 4096,), dtype=torch.uint8, device=dev)
    dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    gb = y.numel() * 2 / 1e9
    r = {}
    for name, zz, mm in [("none", None, None), ("z", z, None), ("mask", z, mask)]:
        t_rd = timeit(lambda: native.call("picd_bn_bwd_reduce_mask", dz, y, zz, mm, coef[0], coef[1], coef[2], coef[3], 1, sums, P, C))
        t_ba = timeit(lambda: native.call("picd_bn_bwd_apply_mask", dz, y, zz, mm, coef[0], coef[1], coef[2], coef[3], 1, sums, 16, 1.0 / P,
                                          sums, 16, dg, db, out, dres if zz is not None else None, P, C))
        r[name] = (t_rd, t_ba)
    print(H, C, f"{gb:.3f} GB", {k: (round(v[0], 4), round(v[1], 4)) for k, v in r.items()})
