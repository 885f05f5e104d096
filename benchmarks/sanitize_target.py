#!/usr/bin/env python
"""Small-shape pass over every kernel family, meant to run under compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python benchmarks/sanitize_target.py
One forward + backward + fused update of a tiny ResNet-18, an Inception-style concat/pool block and a ViT block."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200 import ops  # noqa: E402
from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss  # noqa: E402
from picd_b200.models import Classifier  # noqa: E402
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
which = sys.argv[1:] or ["resnet18", "vit_tiny", "ops"]
if "resnet18" in which:
    m = FusedDataParallel(Classifier("resnet18", 7, image_size=64).to(dev), OptimizerSpec("sgd", lr=0.01, momentum=0.9))
    m.train()
    x, y = ops.synthetic_batch(4, 3, 64, 64, 7, 0, 0, dev, torch.bfloat16)
    for _ in range(2):
        loss, _ = compute_loss(m(x), y, WeightedCrossEntropy(None))
        loss.backward()
        m.optimizer.step()
    print("resnet18 ok", float(loss))
if "vit_tiny" in which:
    from picd_b200.models.vit import VisionTransformer
    vit = VisionTransformer(image_size=32, patch=16, depth=1, heads=2, dim=128, hidden=256, num_classes=7).to(dev)
    m = FusedDataParallel(vit, OptimizerSpec("adamw", lr=1e-3))
    m.train()
    x, y = ops.synthetic_batch(4, 3, 32, 32, 7, 0, 1, dev, torch.bfloat16)
    loss = ops.cross_entropy(m(x).float(), y)
    loss.backward()
    m.optimizer.step()
    print("vit ok", float(loss))
if "ops" in which:
    a = torch.randn(2, 64, 17, 17, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_()
    b = ops.max_pool2d(a, 3, 2, 0)
    c = ops.avg_pool2d(a, 3, 1, 1)
    d = ops.concat_channels([c, c])
    (b.float().sum() + d.float().sum()).backward()
    w = (torch.randn(64, 1, 5, 5, device=dev) / 5).requires_grad_()
    from picd_b200.ops import native
    e = native.depthwise_conv2d(a, w, 2, (1, 1, 2, 2))
    e.float().sum().backward()
    print("ops ok")
torch.cuda.synchronize()
