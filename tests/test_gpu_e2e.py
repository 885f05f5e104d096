"""End-to-end parity on a GPU (last row): N training steps of the native path against the same
model/seed/batches on the library-op path (PICD_FORCE_TORCH=1); the loss curves must agree within bf16 noise, and
the checkpoint written by one path must load into the other."""
import os

import pytest

pytestmark = pytest.mark.gpu


def _train(force_torch: bool, steps: int = 8, name: str = "resnet18"):
    import torch
    from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss
    from picd_b200.models import Classifier
    from picd_b200.ops import reference as ref
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    os.environ["PICD_FORCE_TORCH"] = "1" if force_torch else "0"
    try:
        dev = torch.device("cuda", 0)
        torch.manual_seed(0)
        model = Classifier(name, 7, image_size=64).to(dev)
        dp = FusedDataParallel(model, OptimizerSpec("sgd", lr=0.02, momentum=0.9, weight_decay=1e-4))
        dp.train()
        loss_fn = WeightedCrossEntropy(torch.tensor([3, 3, 10, 1, 4, 4, 5.0], device=dev))
        losses = []
        for s in range(steps):
            img, lab = ref.synthetic_batch(32, 3, 64, 64, 7, 0, s % 2, "cpu", torch.float32)   # two alternating batches
            out = dp(img.to(dev))
            loss, _ = compute_loss(out, lab.to(dev), loss_fn)
            dp.optimizer.zero_grad()
            loss.backward()
            dp.optimizer.step()
            losses.append(float(loss.item()))
        return losses, dp
    finally:
        os.environ["PICD_FORCE_TORCH"] = "0"


def test_loss_curve_matches_library_path(tmp_path):
    import torch
    from picd_b200.utils import load_checkpoint, save_checkpoint
    native, dp_native = _train(False)
    library, dp_lib = _train(True)
    assert all(l == l for l in native + library)
    assert native[-1] < native[0] and library[-1] < library[0]                 # both learn the two batches
    for a, b in zip(native, library):
        assert abs(a - b) < 0.08 * max(1.0, abs(b)), (native, library)
    # cross-path checkpoint compatibility (same key layout: module.encoder.*)
    save_checkpoint(dp_native, 3, 42.0, str(tmp_path), "resnet18")
    epoch, best, _ = load_checkpoint(dp_lib, str(tmp_path), "resnet18")
    assert (epoch, best) == (3, 42.0)
    a = dp_native.module.encoder.layer2[0].conv1.weight
    b = dp_lib.module.encoder.layer2[0].conv1.weight
    assert torch.equal(a, b)


def test_eval_mode_and_validation_loop():
    """val_epoch on the native path: running statistics are used, accuracy is a percentage, no grad is recorded."""
    import torch
    from picd_b200.data import SyntheticDeviceLoader
    from picd_b200.engine.trainer import val_epoch
    _, dp = _train(False, steps=2)
    dev = torch.device("cuda", 0)
    loader = SyntheticDeviceLoader(16, 64, 7, 2, dev, seed=3)
    acc = val_epoch(0, dp, loader, dev, printer=lambda s: None)
    assert 0.0 <= acc <= 100.0
