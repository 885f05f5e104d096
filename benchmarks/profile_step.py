#!/usr/bin/env python
"""Per-kernel device-time shares of one training step (torch.profiler / CUPTI; shares only, never a bench number).

    python benchmarks/profile_step.py --model resnet50 --batch 256 [--impl torch]
Writes gpurun_out/step_profile_<impl>.txt."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--size", type=int, default=224)
ap.add_argument("--impl", default="native", choices=["native", "torch"])
ap.add_argument("--steps", type=int, default=3)
args = ap.parse_args()
if args.impl == "torch":
    os.environ["PICD_FORCE_TORCH"] = "1"

import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss  # noqa: E402
from picd_b200.models import Classifier  # noqa: E402
from picd_b200.ops import synthetic_batch  # noqa: E402
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec  # noqa: E402

dev = torch.device("cuda")
torch.manual_seed(0)
m = Classifier(args.model, 7, image_size=args.size).to(dev)
dp = FusedDataParallel(m, OptimizerSpec("sgd", lr=0.01, momentum=0.9))
dp.train()
img, lab = synthetic_batch(args.batch, 3, args.size, args.size, 7, 0, 0, dev, torch.bfloat16)
loss_fn = WeightedCrossEntropy(None)


def step():
    loss, _ = compute_loss(dp(img), lab, loss_fn)
    loss.backward()
    dp.optimizer.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages():
    dt = getattr(e, "device_time_total", 0) or getattr(e, "cuda_time_total", 0)
    if dt and e.device_type.name == "CUDA":
        rows.append((dt / args.steps / 1e3, e.count / args.steps, e.key))
rows.sort(reverse=True)
total = sum(r[0] for r in rows)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
path = os.path.join(ROOT, "gpurun_out", f"step_profile_{args.impl}_{args.model}_b{args.batch}.txt")
with open(path, "w") as f:
    f.write(f"# {args.model} batch {args.batch} {args.size}px impl={args.impl}: device time per step {total:.2f} ms "
            f"(sum over kernels, torch.profiler; shares only)\n")
    f.write(f"{'ms/step':>10} {'share':>7} {'calls':>7}  kernel\n")
    for ms, cnt, name in rows[:40]:
        f.write(f"{ms:10.3f} {100 * ms / total:6.1f}% {cnt:7.1f}  {name[:150]}\n")
print(open(path).read())
