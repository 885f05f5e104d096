"""AverageMeter / Accuracy parity with reference utils.py:5-27."""
import torch

from picd_b200.utils import Accuracy, AverageMeter


def test_average_meter_matches_reference_arithmetic():
    m = AverageMeter()
    m.update(2.0, 4)
    m.update(4.0, 4)
    assert m.val == 4.0 and m.sum == 24.0 and m.count == 8 and m.avg == 3.0
    m.reset()
    assert m.count == 0 and m.avg == 0


def test_average_meter_keeps_tensors_lazy():
    m = AverageMeter()
    m.update(torch.tensor(1.0), 2)
    m.update(torch.tensor(3.0), 2)
    assert torch.is_tensor(m.avg)
    val, avg = m.values()
    assert val == 3.0 and avg == 2.0


def test_accuracy_is_int64_vector_like_reference():
    out = torch.tensor([[0.1, 0.9], [0.8, 0.2], [0.3, 0.7]])
    hits = Accuracy(out, torch.tensor([1, 1, 1]))
    assert hits.dtype == torch.int64 and hits.tolist() == [1, 0, 1]
    single = Accuracy(out[0].unsqueeze(0), torch.tensor(1))      # the reference's per-sample call shape
    assert single.tolist() == [1]


def test_top_level_shim_modules():
    import ddp_utils
    import utils
    from dp.loader import ImageDataset  # noqa: F401
    from nn.classifier import Classifier  # noqa: F401
    assert utils.AverageMeter is AverageMeter
    assert callable(ddp_utils.all_gather) and callable(ddp_utils.reduce_tensor) and callable(ddp_utils.barrier)
