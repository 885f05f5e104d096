import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
import torch
from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss
from picd_b200.models import Classifier
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
from picd_b200.utils.meters import Accuracy
dev = torch.device("cuda")
def run(force, lr, amp, steps=150, batch=64, size=64, nc=7):
    os.environ["PICD_FORCE_TORCH"] = "1" if force else "0"
    torch.manual_seed(0)
    gen = torch.Generator(device="cpu").manual_seed(7)
    templates = torch.randn(nc, 3, size, size, generator=gen).to(dev)
    def make(n, seed):
        gg = torch.Generator(device="cpu").manual_seed(seed)
        lab = torch.randint(0, nc, (n,), generator=gg).to(dev)
        img = amp * templates[lab] + torch.randn(n, 3, size, size, generator=gg).to(dev)
        return img.contiguous(memory_format=torch.channels_last), lab
    model = Classifier("resnet18", nc, image_size=size).to(dev)
    dp = FusedDataParallel(model, OptimizerSpec("sgd", lr=lr, momentum=0.9, weight_decay=1e-4))
    loss_fn = WeightedCrossEntropy(None)
    dp.train(); curve = []
    for s in range(steps):
        img, lab = make(batch, 1000 + s)
        loss, _ = compute_loss(dp(img), lab, loss_fn)
        dp.optimizer.zero_grad(); loss.backward(); dp.optimizer.step()
        if s % 15 == 0: curve.append(round(float(loss.item()), 3))
    dp.eval(); hits = []
    with torch.no_grad():
        for s in range(8):
            img, lab = make(batch, 5000 + s)
            hits.append(Accuracy(dp(img).float(), lab))
    os.environ["PICD_FORCE_TORCH"] = "0"
    return float(torch.cat(hits).float().mean().item() * 100), curve
for lr, amp in [(0.05, 0.6), (0.02, 1.0), (0.01, 1.0)]:
    for force in (False, True):
        print("force_torch", force, "lr", lr, "amp", amp, run(force, lr, amp))
