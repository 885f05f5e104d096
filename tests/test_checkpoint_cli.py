"""Checkpoint layout (train.py:131-153,173-188) and CLI flag compatibility (train.py:27-31)."""
import os

import torch

from picd_b200.engine.cli import build_parser
from picd_b200.models import Classifier
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
from picd_b200.utils import BEST, LATEST, checkpoint_path, load_checkpoint, save_checkpoint


def test_cli_accepts_all_local_rank_spellings(monkeypatch):
    p = build_parser()
    assert p.parse_args(["--local_rank", "3", "--datadir", "d"]).local_rank == 3
    assert p.parse_args(["--local-rank=2", "--datadir", "d"]).local_rank == 2
    a = p.parse_args(["--datadir", "x"])
    assert a.batchsize == 4 and a.lr == 0.5e-5 and a.epochs == 100 and a.model == "inceptionv3"
    from picd_b200.parallel.dist import get_local_rank
    monkeypatch.setenv("LOCAL_RANK", "5")
    assert get_local_rank(1) == 5
    monkeypatch.delenv("LOCAL_RANK")
    assert get_local_rank(1) == 1


def test_checkpoint_layout_and_roundtrip(tmp_path):
    torch.manual_seed(0)
    model = FusedDataParallel(Classifier("resnet18", 5, small_input=True), OptimizerSpec("sgd", lr=0.1))
    path = save_checkpoint(model, 7, 81.5, str(tmp_path), "resnet18", BEST)
    assert path == checkpoint_path(str(tmp_path), "resnet18", BEST) and os.path.exists(path)
    blob = torch.load(path, map_location="cpu", weights_only=False)
    assert set(blob) == {"epoch", "best_score", "state_dict"}
    keys = list(blob["state_dict"])
    assert all(k.startswith("module.encoder.") for k in keys)
    assert "module.encoder.conv1.weight" in keys and "module.encoder.fc.6.bias" in keys
    assert "module.encoder.bn1.num_batches_tracked" in keys
    # state_dict() of the wrapper carries the same prefix as the reference's DDP wrapper
    assert set(model.state_dict()) == set(keys)

    torch.manual_seed(1)
    other = FusedDataParallel(Classifier("resnet18", 5, small_input=True), OptimizerSpec("sgd", lr=0.1))
    epoch, best, _ = load_checkpoint(other, str(tmp_path), "resnet18", BEST)
    assert (epoch, best) == (7, 81.5)
    for (k, a), (_, b) in zip(model.state_dict().items(), other.state_dict().items()):
        assert torch.equal(a, b), k
    # parameters are still views of the flat arena after loading
    p = next(other.module.parameters())
    assert p.data_ptr() >= other.arena.master.data_ptr()


def test_partial_key_resume_and_missing_file(tmp_path):
    model = Classifier("resnet18", 5, small_input=True)
    assert load_checkpoint(model, str(tmp_path), "resnet18", BEST) == (0, 0.0, {})
    save_checkpoint(model, 1, 10.0, str(tmp_path), "resnet18", LATEST)
    bigger_head = Classifier("resnet18", 9, small_input=True)     # last Linear has a different shape
    before = bigger_head.encoder.fc[6].weight.clone()
    e, b, _ = load_checkpoint(bigger_head, str(tmp_path), "resnet18", LATEST)
    assert (e, b) == (1, 10.0)
    assert torch.equal(bigger_head.encoder.fc[6].weight, before)              # mismatching key skipped silently
    assert torch.equal(bigger_head.encoder.conv1.weight, model.encoder.conv1.weight)


def test_reference_style_checkpoint_loads():
    """A torchvision-keyed DDP checkpoint (what the reference writes) loads into our Classifier."""
    import torchvision
    tv = torchvision.models.resnet18()
    ours = Classifier("resnet18", 7)
    sd = {"module.encoder." + k: v for k, v in tv.state_dict().items() if not k.startswith("fc.")}
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "resnet18"))
        torch.save({"epoch": 3, "best_score": 50.0, "state_dict": sd}, os.path.join(d, "resnet18", BEST))
        load_checkpoint(ours, d, "resnet18", BEST)
    assert torch.equal(ours.encoder.layer1[0].conv1.weight, tv.layer1[0].conv1.weight)
