#!/usr/bin/env python
"""GPU self-test + micro-benchmarks for libpicd_b200.so.

Each group runs in its own subprocess under a timeout (a trapped kernel poisons only that process) and appends
JSON lines to ``gpurun_out/selftest.jsonl``.  Usage on a GPU box:

    python benchmarks/gpu_selftest.py                 # all groups
    python benchmarks/gpu_selftest.py --group conv    # one group, in-process
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT_DIR = os.path.join(ROOT, "gpurun_out")
GROUPS = ["basic", "gemm", "attn", "conv", "compose", "converge", "model", "perf"]


def emit(rec):
    os.makedirs(OUT_DIR, exist_ok=True)
    with open(os.path.join(OUT_DIR, "selftest.jsonl"), "a") as f:
        f.write(json.dumps(rec) + "\n")
    print(json.dumps(rec), flush=True)


def emit_ok(rec) -> bool:
    """emit() a record that carries its own verdict and hand that verdict back, so callers can `ok &= emit_ok(...)`."""
    emit(rec)
    return bool(rec.get("ok", False))


def relerr(a, b):
    import torch
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-6))


def check(name, got, want, tol, group, extra=None):
    e = relerr(got, want)
    rec = {"group": group, "test": name, "relerr": e, "tol": tol, "ok": bool(e <= tol) and e == e}
    if extra:
        rec.update(extra)
    emit(rec)
    return rec["ok"]


def check_sparse(name, got, want, tol, group, max_bad_frac=2e-5):
    """Like check(), but tolerates a handful of outliers: a ReLU mask flips when a pre-activation lies within rounding
    distance of zero, which happens for a few elements per million and changes that element's gradient completely."""
    got, want = got.float(), want.float()
    err = (got - want).abs() / (want.abs().max() + 1e-6)
    bad = float((err > tol).float().mean())
    rec = {"group": group, "test": name, "relerr_p99.99": float(err.flatten().kthvalue(max(1, int(err.numel() * 0.9999))).values),
           "bad_frac": bad, "tol": tol, "ok": bad <= max_bad_frac}
    emit(rec)
    return rec["ok"]


def time_cuda(fn, iters=10, warmup=3):
    import torch
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


# ------------------------------------------------------------------------------------------ basic
def group_basic():
    import torch
    import torch.nn.functional as F
    from picd_b200 import ops
    from picd_b200.ops import native, reference as ref
    dev = torch.device("cuda")
    ok = True
    g = "basic"
    # synthetic + normalize
    img, lab = ops.synthetic_batch(4, 3, 32, 32, 7, 3, 5, dev, torch.float32)
    rimg, rlab = ref.synthetic_batch(4, 3, 32, 32, 7, 3, 5, "cpu", torch.float32)
    ok &= check("synthetic_fp32", img.cpu(), rimg, 1e-6, g)
    ok &= emit_ok({"group": g, "test": "synthetic_labels", "ok": bool((lab.cpu() == rlab).all())})
    u8 = torch.randint(0, 256, (2, 16, 16, 3), dtype=torch.uint8, device=dev)
    ok &= check("normalize_u8", ops.normalize_u8(u8, torch.bfloat16), ref.normalize_u8(u8), 1e-2, g)

    # channel concat / split (inception blocks) and batched top-1
    parts = [torch.randn(3, c, 9, 7, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
             for c in (64, 96, 8, 320)]
    wide = ops.concat_channels(parts)
    want = torch.cat([q.detach() for q in parts], 1)
    ok &= emit_ok({"group": g, "test": "concat_nhwc_fwd", "ok": bool(torch.equal(wide, want)) and wide.is_contiguous(memory_format=torch.channels_last)})
    dw_ = torch.randn_like(wide)
    wide.backward(dw_)
    offs = [0, 64, 160, 168, 488]
    ok &= emit_ok({"group": g, "test": "concat_nhwc_bwd",
          "ok": all(bool(torch.equal(q.grad, dw_[:, offs[i]:offs[i + 1]])) for i, q in enumerate(parts))})
    for dt in (torch.float32, torch.bfloat16):
        lg = torch.randn(1000, 7, device=dev).to(dt)
        lb = torch.randint(0, 7, (1000,), device=dev)
        hits = ops.top1_hits(lg, lb)
        ok &= emit_ok({"group": g, "test": f"top1_hits[{dt}]", "ok": bool(torch.equal(hits, 1 * (lg.argmax(1) == lb))) and hits.dtype == torch.int64})

    # BN family
    for (N, C, H, W, res, act) in [(8, 64, 14, 14, False, "relu"), (4, 256, 7, 7, True, "relu"), (4, 96, 9, 9, False, "none"),
                                   (2, 2048, 7, 7, True, "relu"), (4, 40, 6, 6, False, "swish"),
                                   # >= 4 MB: these take the TMA-streamed kernels (bn_stream.cu)
                                   (32, 64, 56, 56, False, "relu"), (16, 256, 28, 28, True, "relu"),
                                   (64, 128, 16, 17, False, "none"), (24, 512, 14, 14, True, "relu"),
                                   # channel counts that are not powers of two (inception / efficientnet) on the streamed kernels
                                   (32, 144, 38, 38, False, "swish"), (16, 768, 17, 17, True, "relu"), (48, 192, 35, 35, False, "relu")]:
        x = torch.randn(N, C, H, W, device=dev).mul_(1.5).add_(0.3).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        r = torch.randn(N, C, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last) if res else None
        gamma = (torch.rand(C, device=dev) + 0.5).requires_grad_()
        beta = (torch.randn(C, device=dev) * 0.1).requires_grad_()
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        xn = x.clone().requires_grad_()
        rn = r.clone().requires_grad_() if res else None
        z = native.bn_act(xn, gamma, beta, rm, rv, training=True, momentum=0.1, eps=1e-5, act=act, residual=rn, sync_group=False)
        dz = torch.randn_like(z)
        z.backward(dz)
        xr = x.float().clone().requires_grad_()
        rr = r.float().clone().requires_grad_() if res else None
        g2, b2 = gamma.detach().clone().requires_grad_(), beta.detach().clone().requires_grad_()
        rm2, rv2 = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        zr = ref.batch_norm(xr, g2, b2, rm2, rv2, True, 0.1, 1e-5)
        if res:
            zr = zr + rr
        zr = ref.activation(zr, act)
        zr.backward(dz.float())
        tag = f"bn[{N}x{C}x{H}x{W},res={res},{act}]"
        ok &= check(tag + ".fwd", z, zr, 2e-2, g)
        ok &= check_sparse(tag + ".dx", xn.grad, xr.grad, 3e-2, g)
        ok &= check(tag + ".dgamma", gamma.grad, g2.grad, 3e-2, g)
        ok &= check(tag + ".dbeta", beta.grad, b2.grad, 3e-2, g)
        ok &= check(tag + ".running_var", rv, rv2, 1e-2, g)
        if res:
            ok &= check_sparse(tag + ".dres", rn.grad, rr.grad, 3e-2, g)

    # depthwise convolution (EfficientNet), symmetric and TF-"same" asymmetric padding
    for (N, C, H, K, stride, pad) in [(2, 32, 17, 3, 1, (1, 1, 1, 1)), (2, 96, 16, 3, 2, (0, 0, 1, 1)), (2, 144, 15, 5, 2, (2, 2, 2, 2)),
                                      (3, 40, 9, 5, 1, (2, 2, 2, 2)), (2, 288, 19, 5, 2, (1, 1, 2, 2)), (2, 8, 33, 7, 1, (3, 3, 3, 3)),
                                      (4, 528, 10, 3, 1, (1, 1, 1, 1)), (2, 48, 21, 5, 1, (1, 2, 3, 2)), (1, 16, 6, 3, 1, (0, 1, 2, 1))]:
        x = torch.randn(N, C, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(C, 1, K, K, device=dev) / K).requires_grad_()
        xn = x.clone().requires_grad_()
        y = native.depthwise_conv2d(xn, w, stride, pad)
        dy = torch.randn_like(y)
        y.backward(dy)
        xr = x.float().contiguous().requires_grad_()
        wr = w.detach().clone().requires_grad_()
        pt, pl, pb, pr = pad
        yr = F.conv2d(F.pad(xr, [pl, pr, pt, pb]), wr, None, stride, 0, 1, C)
        yr.backward(dy.float().contiguous())
        tag = f"depthwise[{N}x{C}x{H},k{K},s{stride},pad{pad}]"
        ok &= check(tag + ".fwd", y, yr, 2e-2, g)
        ok &= check(tag + ".dx", xn.grad, xr.grad, 2e-2, g)
        ok &= check(tag + ".dw", w.grad, wr.grad, 2e-2, g)

    # pooling
    for (k, s, p, mode) in [(3, 2, 1, "max"), (3, 2, 0, "max"), (3, 1, 1, "avg"), (5, 3, 0, "avg")]:
        x = torch.randn(2, 64, 17, 17, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        xn = x.clone().requires_grad_()
        y = native.pool2d(xn, k, s, p, mode)
        dy = torch.randn_like(y)
        y.backward(dy)
        # NCHW-contiguous oracle: torch 2.11's CUDA avg_pool2d backward returns wrong values for channels_last
        # fp32 inputs at stride 1 (checked against a box-filter formulation, benchmarks/debug_pool.py)
        xr = x.float().contiguous().requires_grad_()
        yr = (F.max_pool2d if mode == "max" else F.avg_pool2d)(xr, k, s, p)
        yr.backward(dy.float().contiguous())
        ok &= check(f"pool[{mode},{k},{s},{p}].fwd", y, yr, 1e-2, g)
        ok &= check(f"pool[{mode},{k},{s},{p}].bwd", xn.grad, xr.grad, 2e-2, g)
    x = torch.randn(4, 128, 7, 7, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xn = x.clone().requires_grad_()
    y = native.global_avg_pool(xn)
    y.backward(torch.ones_like(y))
    ok &= check("gap.fwd", y, x.float().mean((2, 3)), 1e-2, g)
    ok &= check("gap.bwd", xn.grad, torch.full_like(x, 1 / 49.0), 1e-2, g)
    for (N, C, H, W) in [(3, 144, 40, 40), (2, 32, 75, 75), (5, 8, 33, 31)]:       # pixel-split variant (few samples x channels)
        x = (torch.randn(N, C, H, W, device=dev) + 0.5).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        ok &= check(f"gap_split[{N},{C},{H},{W}].fwd", native.global_avg_pool(x), x.float().mean((2, 3)), 1e-2, g)

    # cross entropy
    lg = torch.randn(37, 7, device=dev, requires_grad=True)
    lb = torch.randint(0, 7, (37,), device=dev)
    w = torch.tensor([3, 3, 10, 1, 4, 4, 5.0], device=dev)
    l = native.cross_entropy(lg, lb, w)
    l.backward()
    lg2 = lg.detach().clone().requires_grad_()
    l2 = F.cross_entropy(lg2, lb, w)
    l2.backward()
    ok &= check("xent.loss", l.reshape(1), l2.reshape(1), 1e-5, g)
    ok &= check("xent.grad", lg.grad, lg2.grad, 1e-4, g)

    # layernorm
    x = torch.randn(50, 768, device=dev).to(torch.bfloat16)
    gm, bt = (torch.rand(768, device=dev) + 0.5).requires_grad_(), torch.randn(768, device=dev).requires_grad_()
    xn = x.clone().requires_grad_()
    y = native.layer_norm(xn, gm, bt, 1e-6)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr, gm2, bt2 = x.float().requires_grad_(), gm.detach().clone().requires_grad_(), bt.detach().clone().requires_grad_()
    yr = F.layer_norm(xr, (768,), gm2, bt2, 1e-6)
    yr.backward(dy.float())
    ok &= check("layernorm.fwd", y, yr, 2e-2, g)
    ok &= check("layernorm.dx", xn.grad, xr.grad, 3e-2, g)
    ok &= check("layernorm.dgamma", gm.grad, gm2.grad, 3e-2, g)
    ok &= check("layernorm.dbeta", bt.grad, bt2.grad, 3e-2, g)

    # residual add fused into LayerNorm (transformer blocks): (s, y) = (x + d, LN(x + d)); both outputs carry gradient
    for (M, D) in [(50, 768), (333, 256), (64, 1024)]:
        x = torch.randn(M, D, device=dev).to(torch.bfloat16)
        d = torch.randn(M, D, device=dev).to(torch.bfloat16)
        gm, bt = (torch.rand(D, device=dev) + 0.5).requires_grad_(), torch.randn(D, device=dev).requires_grad_()
        xn, dn = x.clone().requires_grad_(), d.clone().requires_grad_()
        sn, yn = ops.add_layer_norm(xn, dn, gm, bt, 1e-6)
        gs, gy = torch.randn_like(sn), torch.randn_like(yn)
        torch.autograd.backward([sn, yn], [gs, gy])
        xr, dr = x.float().requires_grad_(), d.float().requires_grad_()
        gm2, bt2 = gm.detach().clone().requires_grad_(), bt.detach().clone().requires_grad_()
        sr = (xr + dr).to(torch.bfloat16).float() + 0 * (xr + dr)       # the kernel normalises the bf16-rounded sum
        sr = xr + dr
        yr = F.layer_norm(sr, (D,), gm2, bt2, 1e-6)
        torch.autograd.backward([sr, yr], [gs.float(), gy.float()])
        ok &= check(f"add_layernorm[{M},{D}].sum", sn, sr, 1e-2, g)
        ok &= check(f"add_layernorm[{M},{D}].y", yn, yr, 2e-2, g)
        ok &= check(f"add_layernorm[{M},{D}].dx", xn.grad, xr.grad, 3e-2, g)
        ok &= check(f"add_layernorm[{M},{D}].ddelta", dn.grad, dr.grad, 3e-2, g)
        ok &= check(f"add_layernorm[{M},{D}].dgamma", gm.grad, gm2.grad, 3e-2, g)
        ok &= check(f"add_layernorm[{M},{D}].dbeta", bt.grad, bt2.grad, 3e-2, g)

    # small linear
    # the last three take the tcgen05 GEMM path with the separate activation kernels (GELU: fast erf; dbias fused in act_bwd)
    for (M, K, N, act) in [(16, 2048, 128, "relu"), (16, 32, 7, None), (5, 64, 32, "relu"), (1024, 768, 3072, "gelu"),
                           (788, 256, 512, "swish"), (512, 768, 768, None)]:
        x = torch.randn(M, K, device=dev).to(torch.bfloat16)
        wt = (torch.randn(N, K, device=dev) / K ** 0.5).requires_grad_()
        b = torch.randn(N, device=dev).requires_grad_()
        xn = x.clone().requires_grad_()
        y = native.linear(xn, wt, b, act)
        dy = torch.randn_like(y)
        y.backward(dy)
        xr, w2, b2 = x.float().requires_grad_(), wt.detach().clone().requires_grad_(), b.detach().clone().requires_grad_()
        yr = ref.linear(xr, w2, b2, act)
        yr.backward(dy.float())
        ok &= check(f"small_linear[{M},{K},{N},{act}].fwd", y, yr, 2e-2, g)
        ok &= check(f"small_linear[{M},{K},{N},{act}].dx", xn.grad, xr.grad, 3e-2, g)
        ok &= check(f"small_linear[{M},{K},{N},{act}].dw", wt.grad, w2.grad, 3e-2, g)
        ok &= check(f"small_linear[{M},{K},{N},{act}].db", b.grad, b2.grad, 3e-2, g)

    # fused update, single GPU
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    for kind in ["sgd", "adam", "adamw"]:
        torch.manual_seed(0)
        lin = torch.nn.Linear(96, 64).to(dev)
        refm = torch.nn.Linear(96, 64).to(dev)
        refm.load_state_dict(lin.state_dict())
        spec = OptimizerSpec(kind=kind, lr=0.1 if kind == "sgd" else 1e-2, momentum=0.9, weight_decay=0.01)
        dp = FusedDataParallel(lin, spec)
        opt = {"sgd": lambda p: torch.optim.SGD(p, lr=0.1, momentum=0.9, weight_decay=0.01),
               "adam": lambda p: torch.optim.Adam(p, lr=1e-2, weight_decay=0.01),
               "adamw": lambda p: torch.optim.AdamW(p, lr=1e-2, weight_decay=0.01)}[kind](refm.parameters())
        for it in range(3):
            x = torch.randn(8, 96, device=dev)
            loss = dp(x).square().mean()
            loss.backward()
            dp.optimizer.step()
            opt.zero_grad()
            refm(x).square().mean().backward()
            opt.step()
        ok &= check(f"fused_update[{kind}].weight", lin.weight, refm.weight, 1e-4, g)
        ok &= check(f"fused_update[{kind}].shadow", lin.weight._shadow, refm.weight, 1e-2, g)
        ok &= emit_ok({"group": g, "test": f"fused_update[{kind}].grad_zeroed", "ok": bool(dp.arena.grad.abs().max().item() == 0)})
    # squeeze-excite gate (+ stochastic-depth mask): out = x * s[n, c] * m[n]
    for (N, C, H, W, masked) in [(4, 144, 19, 19, False), (3, 32, 75, 75, True), (5, 1392, 10, 10, True), (2, 8, 7, 9, False)]:
        torch.manual_seed(N + C)
        x = torch.randn(N, C, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        sg = torch.rand(N, C, device=dev).to(torch.bfloat16)
        m = ((torch.rand(N, device=dev) < 0.7).float() / 0.7) if masked else None
        xn, sn = x.clone().requires_grad_(), sg.clone().requires_grad_()
        out = ops.channel_scale(xn, sn, m)
        dy = torch.randn_like(out)
        out.backward(dy)
        xr, sr = x.float().requires_grad_(), sg.float().requires_grad_()
        outr = xr * sr[:, :, None, None] * (1.0 if m is None else m.view(-1, 1, 1, 1))
        outr.backward(dy.float())
        tag = f"channel_scale[{N},{C},{H},{W},mask={masked}]"
        ok &= check(tag + ".fwd", out, outr, 1e-2, g)
        ok &= check(tag + ".dx", xn.grad, xr.grad, 1e-2, g)
        ok &= check(tag + ".ds", sn.grad, sr.grad, 1.5e-2, g)
    return ok


# ------------------------------------------------------------------------------------------ gemm
def group_gemm():
    import torch
    from picd_b200.ops import native
    dev = torch.device("cuda")
    ok = True
    g = "gemm"
    for (M, N, K) in [(128, 64, 64), (256, 128, 128), (1000, 256, 192), (4096, 768, 768), (50432 // 8, 3072, 768), (300, 2304, 768),
                      (2100, 8192, 768)]:      # 32 column tiles do not divide the SM pairs: round-robin tile schedule, odd tail

        a = torch.randn(M, K, device=dev).to(torch.bfloat16)
        b = (torch.randn(N, K, device=dev) / K ** 0.5).to(torch.bfloat16)
        bias = torch.randn(N, device=dev)
        d = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        native.call("picd_gemm_nt", a, b, d, bias, 0, M, N, K)
        torch.cuda.synchronize()
        want = a.float() @ b.float().t() + bias
        ok &= check(f"gemm_nt[{M},{N},{K}]", d, want, 1.5e-2, g)
        d2 = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        native.call("picd_gemm_nt", a, b, d2, None, 1, M, N, K)
        ok &= check(f"gemm_nt_relu[{M},{N},{K}]", d2, (a.float() @ b.float().t()).relu(), 1.5e-2, g)
        # dgrad-style: D[M,K] = A[M,N] B[N,K]
        dy = torch.randn(M, N, device=dev).to(torch.bfloat16)
        dx = torch.empty(M, K, dtype=torch.bfloat16, device=dev)
        native.call("picd_gemm_nn", dy, b, dx, M, N, K)
        ok &= check(f"gemm_nn[{M},{N},{K}]", dx, dy.float() @ b.float(), 1.5e-2, g)
        dw = torch.zeros(N, K, dtype=torch.float32, device=dev)
        native.call("picd_gemm_tn_f32", dy, a, dw, M, N, K)
        ok &= check(f"gemm_tn[{M},{N},{K}]", dw, dy.float().t() @ a.float(), 1.5e-2, g)
        # wgrad + bias gradient (ones-tile MMA) in one kernel
        dw2 = torch.zeros(N, K, dtype=torch.float32, device=dev)
        db = torch.zeros(N, dtype=torch.float32, device=dev)
        native.call("picd_gemm_tn_f32_bias", dy, a, dw2, db, M, N, K)
        ok &= check(f"gemm_tn_bias[{M},{N},{K}].dw", dw2, dy.float().t() @ a.float(), 1.5e-2, g)
        ok &= check(f"gemm_tn_bias[{M},{N},{K}].db", db, dy.float().sum(0), 1e-2, g)
    # transformer MLP chain: Linear+GELU -> Linear.  The second layer's dgrad multiplies by gelu'(pre) in its epilogue
    # (picd_gemm_nn_actgrad) and both bias gradients come out of the wgrad kernels (ones-tile MMA)
    from picd_b200 import ops
    for (M, D, Hd, act) in [(1576, 768, 3072, "gelu"), (640, 256, 512, "swish")]:
        torch.manual_seed(3)
        x = torch.randn(M, D, device=dev).to(torch.bfloat16)
        w1 = (torch.randn(Hd, D, device=dev) / D ** 0.5).requires_grad_()
        b1 = (torch.randn(Hd, device=dev) * 0.1).requires_grad_()
        w2 = (torch.randn(D, Hd, device=dev) / Hd ** 0.5).requires_grad_()
        b2 = (torch.randn(D, device=dev) * 0.1).requires_grad_()
        xn = x.clone().requires_grad_()
        y = ops.linear(ops.linear(xn, w1, b1, act), w2, b2, None)
        dy = torch.randn_like(y)
        y.backward(dy)
        xr = x.float().requires_grad_()
        p2 = [t.detach().clone().requires_grad_() for t in (w1, b1, w2, b2)]
        hr = torch.nn.functional.linear(xr, p2[0].to(torch.bfloat16).float(), p2[1])
        hr = torch.nn.functional.gelu(hr) if act == "gelu" else torch.nn.functional.silu(hr)
        yr = torch.nn.functional.linear(hr, p2[2].to(torch.bfloat16).float(), p2[3])
        yr.backward(dy.float())
        ok &= check(f"mlp[{M},{D},{Hd},{act}].fwd", y, yr, 2e-2, g)
        ok &= check(f"mlp[{M},{D},{Hd},{act}].dx", xn.grad, xr.grad, 3e-2, g)
        for nm, a_, b_ in zip(("dw1", "db1", "dw2", "db2"), (w1, b1, w2, b2), p2):
            ok &= check(f"mlp[{M},{D},{Hd},{act}].{nm}", a_.grad, b_.grad, 3e-2, g)
    # throughput of the big ones
    for (M, N, K) in [(8192, 8192, 8192), (50432, 3072, 768), (50432, 768, 3072)]:
        a = torch.randn(M, K, device=dev).to(torch.bfloat16)
        b = torch.randn(N, K, device=dev).to(torch.bfloat16)
        d = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        ms = time_cuda(lambda: native.call("picd_gemm_nt", a, b, d, None, 0, M, N, K))
        ms_ref = time_cuda(lambda: torch.matmul(a, b.t()))
        ok &= check(f"gemm_nt_perf[{M},{N},{K}].value", d, torch.matmul(a, b.t()), 1.5e-2, g)
        emit({"group": g, "test": f"gemm_nt_perf[{M},{N},{K}]", "ms": ms, "tflops": 2.0 * M * N * K / ms / 1e9,
              "cublas_ms": ms_ref, "cublas_tflops": 2.0 * M * N * K / ms_ref / 1e9, "ok": True})
    return ok


# ------------------------------------------------------------------------------------------ attention
def group_attn():
    import torch
    from picd_b200.ops import native, reference as ref
    dev = torch.device("cuda")
    ok = True
    g = "attn"
    for (B, S, H) in [(2, 197, 3), (3, 64, 2), (2, 128, 4), (1, 256, 2), (5, 130, 1)]:
        torch.manual_seed(S)
        qkv = (torch.randn(B * S, 3 * H * 64, device=dev) * 0.8).to(torch.bfloat16).requires_grad_()
        o = native.attention_packed(qkv, B, S, H)
        do = torch.randn_like(o)
        o.backward(do)
        q32 = qkv.detach().float().requires_grad_()
        t = q32.view(B, S, 3, H, 64).permute(2, 0, 3, 1, 4)
        orf = ref.attention(t[0], t[1], t[2]).permute(0, 2, 1, 3).reshape(B * S, H * 64)
        orf.backward(do.float())
        tag = f"attention[B{B},S{S},H{H}]"
        ok &= check(tag + ".fwd", o, orf, 2e-2, g)
        dq = qkv.grad.float().view(B, S, 3, H * 64)
        dr = q32.grad.view(B, S, 3, H * 64)
        for i, nm in enumerate(("dq", "dk", "dv")):
            ok &= check(f"{tag}.{nm}", dq[:, :, i], dr[:, :, i], 3e-2, g)
    # throughput at ViT-B/16 size
    B, S, H = 256, 197, 12
    qkv = torch.randn(B * S, 3 * H * 64, device=dev).to(torch.bfloat16).requires_grad_()
    o = native.attention_packed(qkv, B, S, H)
    do = torch.randn_like(o)
    ms_f = time_cuda(lambda: native.attention_packed(qkv.detach(), B, S, H))
    def fb():
        qkv.grad = None
        native.attention_packed(qkv, B, S, H).backward(do)
    ms_fb = time_cuda(fb)
    t = qkv.detach().view(B, S, 3, H, 64).permute(2, 0, 3, 1, 4)
    ms_lib = time_cuda(lambda: torch.nn.functional.scaled_dot_product_attention(t[0], t[1], t[2]))
    emit({"group": g, "test": "attention_perf[B256,S197,H12]", "fwd_ms": ms_f, "fwd_bwd_ms": ms_fb, "library_fwd_ms": ms_lib,
          "fwd_tflops": 4.0 * B * H * S * S * 64 / ms_f / 1e9, "ok": True})
    return ok


# ------------------------------------------------------------------------------------------ conv
CONV_CASES = [
    # N, H, W, Cin, Cout, R, S, stride, pad
    (2, 56, 56, 64, 64, 1, 1, 1, 0), (2, 56, 56, 64, 256, 1, 1, 1, 0), (2, 56, 56, 256, 64, 1, 1, 1, 0),
    (2, 56, 56, 64, 64, 3, 3, 1, 1), (2, 56, 56, 128, 128, 3, 3, 2, 1), (2, 56, 56, 256, 512, 1, 1, 2, 0),
    (4, 14, 14, 256, 256, 3, 3, 1, 1), (4, 14, 14, 1024, 2048, 1, 1, 2, 0), (4, 7, 7, 512, 512, 3, 3, 1, 1),
    (4, 7, 7, 2048, 512, 1, 1, 1, 0), (3, 17, 17, 128, 128, 1, 7, 1, (0, 3)), (3, 17, 17, 160, 192, 7, 1, 1, (3, 0)),
    (3, 35, 35, 48, 64, 5, 5, 1, 2), (3, 35, 35, 288, 384, 3, 3, 2, 0), (2, 9, 9, 80, 192, 3, 3, 1, 0),
    (2, 224, 224, 3, 64, 7, 7, 2, 3), (2, 32, 32, 3, 64, 3, 3, 1, 1), (2, 64, 64, 3, 768, 16, 16, 16, 0),
    # 3x3 stride-1 shapes that take the halo path (one activation box per column shift, resident weights for 64->64)
    (2, 28, 28, 128, 128, 3, 3, 1, 1), (3, 30, 37, 64, 96, 3, 3, 1, 1), (2, 56, 56, 64, 192, 3, 3, 1, 1),
    (2, 28, 28, 192, 64, 3, 3, 1, 1), (5, 33, 29, 64, 64, 3, 3, 1, 1),
    # dgrads whose OUTPUT is wide (two 256-column tiles; CTA pairs) -- the fused BN-backward reduction walks both passes
    (4, 14, 14, 512, 512, 3, 3, 1, 1), (8, 28, 28, 512, 128, 1, 1, 1, 0), (4, 28, 28, 512, 512, 3, 3, 2, 1),
    (4, 14, 14, 1024, 256, 1, 1, 1, 0), (3, 7, 7, 2048, 512, 1, 1, 1, 0), (2, 57, 55, 256, 64, 1, 1, 1, 0),
]


def group_conv():
    import torch
    import torch.nn.functional as F
    from picd_b200.ops import native
    dev = torch.device("cuda")
    ok = True
    g = "conv"
    for case in CONV_CASES:
        N, H, W, Cin, Cout, R, S, stride, pad = case
        torch.manual_seed(1)
        x = torch.randn(N, Cin, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(Cout, Cin, R, S, device=dev) / (Cin * R * S) ** 0.5).contiguous(memory_format=torch.channels_last)
        w.requires_grad_()
        first = Cin % 8 != 0
        xn = x.clone()
        if not first:
            xn.requires_grad_()
        y = native.conv2d(xn, w, None, stride, pad)
        dy = torch.randn_like(y)
        y.backward(dy)
        xr = x.float().requires_grad_()
        wr = w.detach().to(torch.bfloat16).float().requires_grad_()
        yr = F.conv2d(xr, wr, None, stride, pad)
        yr.backward(dy.float())
        tag = f"conv{list(case)}"
        ok &= check(tag + ".fprop", y, yr, 1.5e-2, g)
        ok &= check(tag + ".wgrad", w.grad, wr.grad, 2e-2, g)
        if not first:
            ok &= check(tag + ".dgrad", xn.grad, xr.grad, 2e-2, g)
        # fused statistics epilogue
        if not first:
            stats_r = native.stat_buffer(Cout, dev)
            y2, _ = native._conv_forward(x, w.detach(), stride, pad, stats_r, None, None)
            stats = stats_r.view(-1, 2 * Cout).sum(0)
            yf = y2.float()
            ok &= check(tag + ".stats_sum", stats[:Cout], yf.sum((0, 2, 3)), 1e-2, g)
            ok &= check(tag + ".stats_sumsq", stats[Cout:], (yf * yf).sum((0, 2, 3)), 1e-2, g)
        # dgrad with the fused BatchNorm-backward reduction of the upstream layer (picd_conv_dgrad_bnreduce): dx must be
        # bit-identical to the plain dgrad, the sums must match the fp32 evaluation on that dx
        if not first:
            for relu in (True, False):
                gen = torch.Generator(device="cpu").manual_seed(11)
                y_up = (torch.randn(N, Cin, H, W, generator=gen) * 1.5 + 0.3).to(dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
                coef = torch.stack([torch.randn(Cin, generator=gen), torch.randn(Cin, generator=gen) * 0.5,
                                    torch.randn(Cin, generator=gen) * 0.3 + 0.3, torch.rand(Cin, generator=gen) + 0.5]).to(dev)
                rec = native.BnRecord()
                rec.y, rec.coef, rec.relu, rec.eligible = y_up, coef, relu, True
                _, info = native._conv_forward(x, w.detach(), stride, pad, None, None, None)
                dyb = dy.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
                keep = os.environ.get("PICD_FUSE_BN_REDUCE")
                os.environ["PICD_FUSE_BN_REDUCE"] = "all"      # also the shapes the engine leaves to the separate kernel
                try:
                    dx_f, _ = native._conv_backward(info, dyb, w.detach(), True, False, src_bn=rec)
                finally:
                    if keep is None:
                        os.environ.pop("PICD_FUSE_BN_REDUCE")
                    else:
                        os.environ["PICD_FUSE_BN_REDUCE"] = keep
                dx_p, _ = native._conv_backward(info, dyb, w.detach(), True, False)
                same = bool(torch.equal(dx_f, dx_p)) and rec.sums is not None
                ok &= emit_ok({"group": g, "test": tag + f".dgrad_bnreduce[relu={relu}].dx_identical", "ok": same})
                if rec.sums is not None:
                    sums = rec.sums.view(-1, 2 * Cin).sum(0)
                    yf, gf = y_up.float(), dx_p.float()
                    sc, sh, mu, istd = [c.view(1, -1, 1, 1) for c in coef]
                    gm = gf * ((yf * sc + sh) > 0) if relu else gf
                    ok &= check(tag + f".dgrad_bnreduce[relu={relu}].sum_g", sums[:Cin], gm.sum((0, 2, 3)), 1e-2, g)
                    ok &= check(tag + f".dgrad_bnreduce[relu={relu}].sum_g_xhat", sums[Cin:],
                                (gm * (yf - mu) * istd).sum((0, 2, 3)), 1e-2, g)
    ok &= stem_checks()
    return ok


def _stem_run(mode, x, w0, g0, b0, stride, pad, r):
    """max_pool(relu(BN(conv(x)))) and its gradients: 'fused' (csrc/ops/stem.cu), 'unfused' (native conv_bn_act + pool2d) or
    'oracle' (torch, fp32)."""
    import torch
    import torch.nn.functional as F
    from picd_b200.ops import native
    C = w0.shape[0]
    w, gam, bet = w0.clone().requires_grad_(), g0.clone().requires_grad_(), b0.clone().requires_grad_()
    rm, rv = torch.zeros(C, device=x.device), torch.ones(C, device=x.device)
    first = x.shape[1] % 8 != 0
    if mode == "oracle":
        xi = x.float()
        if not first:
            xi.requires_grad_()
        # the same op in fp32 with the two storage roundings (conv output, activation) made explicit as straight-through
        # casts: a max-pool over bf16 values has ties an unrounded oracle would break differently
        def ste(t):
            return t + (t.to(torch.bfloat16).float() - t).detach()
        y = F.conv2d(xi, w.to(torch.bfloat16).float(), None, stride, pad)
        z = ste(F.relu(F.batch_norm(ste(y), rm, rv, gam, bet, True, 0.1, 1e-5)))
        out = F.max_pool2d(z, 3, 2, 1)
    else:
        xi = x.clone()
        if not first:
            xi.requires_grad_()
        kw = dict(stride=stride, padding=pad, training=True, momentum=0.1, eps=1e-5, sync_group=False)
        if mode == "fused":
            out = native.conv_bn_relu_maxpool(xi, w, gam, bet, rm, rv, **kw)
        else:
            z = native.conv_bn_act(xi, w, gam, bet, rm, rv, act="relu", residual=None, **kw)
            out = native.pool2d(z, 3, 2, 1, mode="max")
    out.backward(r.to(out.dtype))
    return {"out": out.detach().float(), "dw": w.grad.float(), "dgamma": gam.grad.float(), "dbeta": bet.grad.float(),
            "dx": None if first else xi.grad.float(), "rm": rm, "rv": rv}


def stem_checks():
    """Fused stem == the two separate native ops (same arg-max rule, so the pooled output is bit-identical and the gradients
    differ only by the bf16 rounding of the dz tensor the fused form never writes) and == the fp32 oracle."""
    import torch
    dev = torch.device("cuda")
    ok = True
    g = "conv"
    #        N  Cin  H    W   Cout R  stride pad
    cases = [(4, 16, 30, 30, 64, 3, 1, 1), (2, 3, 64, 64, 64, 7, 2, 3), (3, 8, 29, 31, 32, 3, 1, 1), (2, 16, 18, 18, 128, 3, 1, 1)]
    for N, Cin, H, W, C, R, stride, pad in cases:
        torch.manual_seed(5)
        x = torch.randn(N, Cin, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w0 = (torch.randn(C, Cin, R, R, device=dev) / (Cin * R * R) ** 0.5).contiguous(memory_format=torch.channels_last)
        g0, b0 = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev) * 0.2
        OH = ((H + 2 * pad - R) // stride + 1 + 2 - 3) // 2 + 1
        OW = ((W + 2 * pad - R) // stride + 1 + 2 - 3) // 2 + 1
        r = torch.randn(N, C, OH, OW, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        res = {m: _stem_run(m, x, w0, g0, b0, stride, pad, r) for m in ("fused", "unfused", "oracle")}
        tag = f"stem{[N, Cin, H, W, C, R, stride, pad]}"
        ok &= emit_ok({"group": g, "test": tag + ".out_identical_to_unfused",
                       "ok": bool(torch.equal(res["fused"]["out"], res["unfused"]["out"]))})
        ok &= check(tag + ".out", res["fused"]["out"], res["oracle"]["out"], 2e-2, g)
        for k in ("dw", "dgamma", "dbeta", "rm", "rv"):
            ok &= check(tag + f".{k}", res["fused"][k], res["oracle"][k], 3e-2, g)
            ok &= check(tag + f".{k}.vs_unfused", res["fused"][k], res["unfused"][k], 1e-2, g)
        # eval mode under no_grad (validation epochs): running statistics, the same single pass
        from picd_b200.ops import native
        rm = torch.randn(C, device=dev) * 0.1
        rv = torch.rand(C, device=dev) + 0.5
        with torch.no_grad():
            kw = dict(stride=stride, padding=pad, training=False, momentum=0.1, eps=1e-5, sync_group=False)
            ev_f = native.conv_bn_relu_maxpool(x, w0, g0, b0, rm.clone(), rv.clone(), **kw)
            ev_u = native.pool2d(native.conv_bn_act(x, w0, g0, b0, rm.clone(), rv.clone(), act="relu", residual=None, **kw), 3, 2, 1,
                                 mode="max")
        ok &= emit_ok({"group": g, "test": tag + ".eval_identical_to_unfused", "ok": bool(torch.equal(ev_f, ev_u))})
        if res["fused"]["dx"] is not None:
            ok &= check_sparse(tag + ".dx", res["fused"]["dx"], res["oracle"]["dx"], 4e-2, g, max_bad_frac=2e-3)
            ok &= check(tag + ".dx.vs_unfused", res["fused"]["dx"], res["unfused"]["dx"], 1.5e-2, g)
    return ok


# ------------------------------------------------------------------------------------------ composition
def _stage_grads(build, x_shape, mode, dev, seed=0):
    """loss = <stage(x), r> for a fixed random r; returns (output, {param: grad}, input grad) in fp32."""
    import torch
    os.environ["PICD_FORCE_TORCH"] = "0" if mode == "native" else "1"
    os.environ["PICD_REF_FP32"] = "1" if mode == "torch_fp32" else "0"
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False      # a real fp32 oracle
    try:
        torch.manual_seed(seed)
        stage = build().to(dev)
        stage.train()
        g = torch.Generator(device="cpu").manual_seed(seed + 1)
        x = torch.randn(*x_shape, generator=g).to(dev)
        r = None
        x = (x.float() if mode == "torch_fp32" else x.to(torch.bfloat16)).contiguous(memory_format=torch.channels_last)
        x.requires_grad_()
        out = stage(x)
        r = torch.randn(out.shape, generator=g).to(dev).to(out.dtype)
        (out.float() * r.float()).mean().backward()
        grads = {k: p.grad.detach().float().clone() for k, p in stage.named_parameters() if p.grad is not None}
        return out.detach().float(), grads, x.grad.detach().float()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
        os.environ["PICD_FORCE_TORCH"] = "0"
        os.environ["PICD_REF_FP32"] = "0"


def group_compose():
    """Composition of the fused paths (residual forks, reduce-add / fused-reduction dgrads, bit-mask BN backward, stride-2
    parity views, n-way inception forks) at a scale where the check can FAIL: a few blocks at batch 128, where the
    bf16 library path agrees with the fp32 oracle to cos >= 0.96 for EVERY parameter (measured 0.973 - 0.988: the noise
    floor of bf16 activations in BatchNorm parameter gradients); the native path must be within 0.015 of that yardstick
    for every parameter and for the input gradient -- a wrong sign, a dropped branch or a wrong mask gives cos < 0.9 (this
    test is what exposed torch's own channels_last avg_pool2d backward, see ops/reference.avg_pool2d).  (The whole-network
    gradient comparison in group_model is dominated by bf16 chaos through 50+ layers -- cos ~0.2 for either path -- and
    only guards the loss and the logits.)"""
    import torch
    import torch.nn as nn
    from picd_b200.models import inception, resnet
    dev = torch.device("cuda")
    ok = True
    g = "compose"

    def resnet_stage():
        # a layer1-style bottleneck with a projection shortcut, an identity-shortcut one, then a stride-2 block
        ds1 = nn.Sequential(resnet.Conv2d(64, 256, 1, 1), resnet.BatchNorm2d(256))
        ds2 = nn.Sequential(resnet.Conv2d(256, 512, 1, 2), resnet.BatchNorm2d(512))
        return nn.Sequential(resnet.Bottleneck(64, 64, 1, ds1), resnet.Bottleneck(256, 64),
                             resnet.Bottleneck(256, 128, 2, ds2))

    def basic_stage():
        ds = nn.Sequential(resnet.Conv2d(64, 128, 1, 2), resnet.BatchNorm2d(128))
        return nn.Sequential(resnet.BasicBlock(64, 64), resnet.BasicBlock(64, 128, 2, ds), resnet.BasicBlock(128, 128))

    def inception_stage():
        return nn.Sequential(inception.InceptionA(192, 32), inception.InceptionA(256, 64))

    cases = [("resnet50_stage", resnet_stage, (128, 64, 32, 32)), ("resnet18_stage", basic_stage, (128, 64, 32, 32)),
             ("inception_stage", inception_stage, (128, 192, 17, 17))]
    cos = lambda a, b: float(torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0))
    for name, build, shape in cases:
        try:
            o1, g1, dx1 = _stage_grads(build, shape, "native", dev)
            o0, g0, dx0 = _stage_grads(build, shape, "torch_bf16", dev)
            o2, g2, dx2 = _stage_grads(build, shape, "torch_fp32", dev)
        except Exception as e:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            emit({"group": g, "test": name, "ok": False, "error": repr(e)[:400]})
            ok = False
            continue
        ok &= check(f"{name}.out", o1, o2, max(0.03, 1.5 * relerr(o0, o2)), g)
        worst_margin, worst_key, yard_min, nat_min = 1.0, None, 1.0, 1.0
        missing = [k for k in g2 if k not in g1]
        for k in list(g2) + ["<input>"]:
            a, b, c = (dx1, dx0, dx2) if k == "<input>" else (g1.get(k), g0[k], g2[k])
            if a is None:
                continue
            cn, ct = cos(a, c), cos(b, c)
            yard_min, nat_min = min(yard_min, ct), min(nat_min, cn)
            if cn - ct < worst_margin:
                worst_margin, worst_key = cn - ct, k
        good = (not missing) and yard_min >= 0.96 and worst_margin >= -0.015
        ok &= emit_ok({"group": g, "test": f"{name}.grads_vs_fp32", "native_min_cos": nat_min, "library_bf16_min_cos": yard_min,
                       "worst_margin": worst_margin, "worst_param": worst_key, "missing": missing[:5], "ok": good})
    return ok


# ------------------------------------------------------------------------------------------ convergence
def _train_to_accuracy(force_torch: bool, dev, steps=150, batch=64, size=64, nc=7):
    """ResNet-18 on a separable device-generated set (class = one of `nc` fixed random templates + noise); returns the
    validation accuracy in percent on fresh noise."""
    import torch
    from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss
    from picd_b200.models import Classifier
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    from picd_b200.utils.meters import Accuracy
    os.environ["PICD_FORCE_TORCH"] = "1" if force_torch else "0"
    try:
        torch.manual_seed(0)
        gen = torch.Generator(device="cpu").manual_seed(7)
        templates = torch.randn(nc, 3, size, size, generator=gen).to(dev)

        def make(n, seed):
            gg = torch.Generator(device="cpu").manual_seed(seed)
            lab = torch.randint(0, nc, (n,), generator=gg).to(dev)
            img = templates[lab] + torch.randn(n, 3, size, size, generator=gg).to(dev)
            return img.contiguous(memory_format=torch.channels_last), lab

        model = Classifier("resnet18", nc, image_size=size).to(dev)
        dp = FusedDataParallel(model, OptimizerSpec("sgd", lr=0.01, momentum=0.9, weight_decay=1e-4))
        loss_fn = WeightedCrossEntropy(None)
        dp.train()
        for s in range(steps):
            img, lab = make(batch, 1000 + s)
            loss, _ = compute_loss(dp(img), lab, loss_fn)
            dp.optimizer.zero_grad()
            loss.backward()
            dp.optimizer.step()
        dp.eval()
        hits = []
        with torch.no_grad():
            for s in range(8):
                img, lab = make(batch, 5000 + s)
                hits.append(Accuracy(dp(img).float(), lab))
        return float(torch.cat(hits).float().mean().item() * 100), float(loss.item())
    finally:
        os.environ["PICD_FORCE_TORCH"] = "0"


def group_converge():
    """Trains to an accuracy: the native path must reach >= 95 % validation accuracy on a separable set and land within
    2 points of the library-op path trained the same way (a sign error in a fused backward path would not)."""
    import torch
    dev = torch.device("cuda")
    acc_n, loss_n = _train_to_accuracy(False, dev)
    acc_t, loss_t = _train_to_accuracy(True, dev)
    return emit_ok({"group": "converge", "test": "resnet18_separable", "native_val_acc": acc_n, "library_val_acc": acc_t,
                    "native_final_loss": loss_n, "library_final_loss": loss_t,
                    "ok": acc_n >= 95.0 and abs(acc_n - acc_t) <= 2.0})


# ------------------------------------------------------------------------------------------ model
def _model_grads(name, mode, batch, size, dev):
    """mode: 'native' (sm_100a kernels, bf16), 'torch_bf16' (library ops, bf16), 'torch_fp32' (the oracle)."""
    import torch
    os.environ["PICD_FORCE_TORCH"] = "0" if mode == "native" else "1"
    os.environ["PICD_REF_FP32"] = "1" if mode == "torch_fp32" else "0"
    from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss
    from picd_b200.models import Classifier
    from picd_b200.ops import reference as ref
    torch.manual_seed(0)
    m = Classifier(name, 7, image_size=size).to(dev)
    m.train()
    for mod in m.modules():           # stochastic layers draw different random numbers on the two paths
        if hasattr(mod, "p") and mod.__class__.__name__ == "Dropout":
            mod.p = 0.0
        if hasattr(mod, "drop_connect"):
            mod.drop_connect = 0.0
    img, lab = ref.synthetic_batch(batch, 3, size, size, 7, 0, 0, "cpu", torch.float32)
    img, lab = img.to(dev), lab.to(dev)
    out = m(img)
    loss, main = compute_loss(out, lab, WeightedCrossEntropy(None))
    loss.backward()
    grads = {k: p.grad.detach().float().clone() for k, p in m.named_parameters() if p.grad is not None}
    os.environ["PICD_FORCE_TORCH"] = "0"
    os.environ["PICD_REF_FP32"] = "0"
    return float(loss.item()), main.detach().float(), grads


def _min_cos(ga, gb):
    import torch
    worst, key = 1.0, None
    for k in gb:
        if k in ga:
            c = float(torch.nn.functional.cosine_similarity(ga[k].flatten(), gb[k].flatten(), dim=0))
            if c < worst:
                worst, key = c, k
    return worst, key


def group_model():
    """Gradients of whole models: the native bf16 path must be as close to the fp32 oracle as the library bf16
    path is (bf16 noise makes early-layer gradients of a random-init deep net only loosely reproducible)."""
    import torch
    dev = torch.device("cuda")
    ok = True
    g = "model"
    for name, batch, size in [("resnet18", 16, 64), ("resnet50", 16, 128), ("inceptionv3", 4, 299), ("vit_b_16", 4, 224),
                              ("efficientnet-b0", 8, 64)]:
        try:
            l1, o1, g1 = _model_grads(name, "native", batch, size, dev)
            l0, o0, g0 = _model_grads(name, "torch_bf16", batch, size, dev)
            l2, o2, g2 = _model_grads(name, "torch_fp32", batch, size, dev)
        except Exception as e:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            emit({"group": g, "test": name, "ok": False, "error": repr(e)[:400]})
            ok = False
            continue
        ok &= emit_ok({"group": g, "test": f"{name}.loss", "native": l1, "torch_bf16": l0, "fp32": l2,
              "ok": abs(l1 - l2) < 0.05 * max(1.0, abs(l2))})
        lib_err = relerr(o0, o2)      # how far the LIBRARY bf16 path is from fp32: the yardstick for bf16 noise
        ok &= check(f"{name}.logits_vs_fp32", o1, o2, max(0.1, 1.5 * lib_err), g, {"torch_bf16_relerr": lib_err})
        missing = [k for k in g2 if k not in g1]
        cn, kn = _min_cos(g1, g2)
        ct, kt = _min_cos(g0, g2)
        mean_cn = sum(float(torch.nn.functional.cosine_similarity(g1[k].flatten(), g2[k].flatten(), dim=0)) for k in g2 if k in g1) / max(1, len(g2))
        mean_ct = sum(float(torch.nn.functional.cosine_similarity(g0[k].flatten(), g2[k].flatten(), dim=0)) for k in g2 if k in g0) / max(1, len(g2))
        # per parameter: wherever the LIBRARY bf16 path still finds the fp32 direction (cosine >= 0.5), the native path must be
        # within 0.4 of it.  Parameters whose bf16 gradient is rounding noise for the library as well (the first blocks' BN biases:
        # cosine around zero either way) carry no signal and made a plain minimum flip between runs.
        cos = torch.nn.functional.cosine_similarity
        worst_gap, worst_param = 0.0, None
        for k in g2:
            if k in g1 and k in g0:
                c0 = float(cos(g0[k].flatten(), g2[k].flatten(), dim=0))
                c1 = float(cos(g1[k].flatten(), g2[k].flatten(), dim=0))
                if c0 >= 0.5 and c0 - c1 > worst_gap:
                    worst_gap, worst_param = c0 - c1, k
        good = (not missing) and mean_cn >= mean_ct - 0.05 and worst_gap <= 0.4     # (a sign error shows up as a gap > 1;
        # the largest gap seen on a correct build is 0.17, on a bias of ResNet-50's last block; stage-level exactness is `compose`)
        emit({"group": g, "test": f"{name}.grads_vs_fp32", "native_min_cos": cn, "native_min_param": kn,
              "torch_bf16_min_cos": ct, "torch_bf16_min_param": kt, "native_mean_cos": mean_cn,
              "torch_bf16_mean_cos": mean_ct, "worst_gap_where_bf16_is_meaningful": worst_gap, "worst_gap_param": worst_param,
              "missing": missing[:5], "ok": good})
        ok &= good
    return ok


# ------------------------------------------------------------------------------------------ perf
def group_perf():
    import torch
    from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss
    from picd_b200.models import Classifier
    from picd_b200.ops import synthetic_batch
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    dev = torch.device("cuda")
    g = "perf"
    for name, batch, size in [("resnet50", 128, 224), ("resnet50", 256, 224)]:
        torch.manual_seed(0)
        m = Classifier(name, 1000 if False else 7, image_size=size).to(dev)
        dp = FusedDataParallel(m, OptimizerSpec(kind="sgd", lr=0.01, momentum=0.9))
        dp.train()
        img, lab = synthetic_batch(batch, 3, size, size, 7, 0, 0, dev, torch.bfloat16)
        loss_fn = WeightedCrossEntropy(None)

        def step():
            out = dp(img)
            loss, _ = compute_loss(out, lab, loss_fn)
            loss.backward()
            dp.optimizer.step()
        ms = time_cuda(step, iters=8, warmup=3)
        emit({"group": g, "test": f"{name}.step[b{batch}]", "ms": ms, "img_per_s": batch / ms * 1e3, "ok": True,
              "mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30})
        del dp, m
        torch.cuda.empty_cache()
    return True


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--group", default=None, choices=GROUPS)
    ap.add_argument("--timeout", type=int, default=420)
    args = ap.parse_args()
    if args.group:
        t0 = time.time()
        try:
            ok = globals()["group_" + args.group]()
        except Exception as e:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            emit({"group": args.group, "test": "EXCEPTION", "ok": False, "error": repr(e)[:600]})
            ok = False
        emit({"group": args.group, "test": "SUMMARY", "ok": bool(ok), "seconds": time.time() - t0})
        sys.exit(0 if ok else 1)
    rc = 0
    for grp in GROUPS:
        try:
            p = subprocess.run([sys.executable, os.path.abspath(__file__), "--group", grp], timeout=args.timeout)
            if p.returncode != 0:
                rc = 1
        except subprocess.TimeoutExpired:
            emit({"group": grp, "test": "TIMEOUT", "ok": False})
            rc = 1
    sys.exit(rc)


if __name__ == "__main__":
    main()
