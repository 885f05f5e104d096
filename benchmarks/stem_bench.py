"""ResNet stem at the headline shape: conv7x7/s2 + BN + ReLU + max-pool 3x3/s2, forward + backward, fused (csrc/ops/stem.cu)
against the two separate native ops.  CUDA events, L2 flushed by the working set itself (411 MB activations)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from picd_b200.ops import native  # noqa: E402


def run(mode, x, w, gam, bet, rm, rv, r, iters=10, warmup=3):
    kw = dict(stride=2, padding=3, training=True, momentum=0.1, eps=1e-5, sync_group=False)
    ts = {"fwd": [], "bwd": []}
    for i in range(warmup + iters):
        for p in (w, gam, bet):
            p.grad = None
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        if mode == "fused":
            out = native.conv_bn_relu_maxpool(x, w, gam, bet, rm, rv, **kw)
        else:
            out = native.pool2d(native.conv_bn_act(x, w, gam, bet, rm, rv, act="relu", residual=None, **kw), 3, 2, 1, mode="max")
        e[1].record()
        out.backward(r)
        e[2].record()
        torch.cuda.synchronize()
        if i >= warmup:
            ts["fwd"].append(e[0].elapsed_time(e[1]))
            ts["bwd"].append(e[1].elapsed_time(e[2]))
    return {k: sorted(v)[len(v) // 2] for k, v in ts.items()}


def main():
    dev = torch.device("cuda")
    B = int(os.environ.get("B", 256))
    torch.manual_seed(0)
    x = torch.randn(B, 3, 224, 224, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(64, 3, 7, 7, device=dev) / 12).contiguous(memory_format=torch.channels_last).requires_grad_()
    gam, bet = torch.ones(64, device=dev, requires_grad=True), torch.zeros(64, device=dev, requires_grad=True)
    rm, rv = torch.zeros(64, device=dev), torch.ones(64, device=dev)
    r = torch.randn(B, 64, 56, 56, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    for mode in ("unfused", "fused"):
        t = run(mode, x, w, gam, bet, rm, rv, r)
        print(json.dumps({"stem": mode, "batch": B, "fwd_ms": round(t["fwd"], 4), "bwd_ms": round(t["bwd"], 4),
                          "total_ms": round(t["fwd"] + t["bwd"], 4)}), flush=True)


if __name__ == "__main__":
    main()
