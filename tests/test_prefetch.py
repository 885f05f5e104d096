"""DevicePrefetcher (data/prefetch.py): order, pass-through of loader attributes, and on a GPU the overlap path
(copy stream + uint8 normalisation) must deliver the same tensors as the synchronous path of the trainer."""
import pytest
import torch

from picd_b200.data import DevicePrefetcher


class _Loader(list):
    sampler = "the-sampler"


def test_prefetcher_preserves_order_and_attributes_on_cpu():
    batches = _Loader({"image": torch.full((2, 3, 4, 4), float(i)), "label": torch.tensor([i, i]), "image_id": ["a", "b"]}
                      for i in range(5))
    pf = DevicePrefetcher(batches, "cpu")
    out = list(pf)
    assert [float(b["image"][0, 0, 0, 0]) for b in out] == [0.0, 1.0, 2.0, 3.0, 4.0]
    assert len(pf) == 5 and pf.sampler == "the-sampler"
    assert list(DevicePrefetcher(_Loader(), "cpu")) == []


@pytest.mark.gpu
def test_prefetcher_matches_synchronous_path_on_gpu():
    from picd_b200 import ops
    from picd_b200.data import SyntheticHostLoader
    dev = torch.device("cuda", 0)
    src = SyntheticHostLoader(8, 32, 7, 6, dev, seed=3)
    got = list(DevicePrefetcher(src.host_iter(6), dev))
    torch.cuda.synchronize()
    assert len(got) == 6
    for s, b in enumerate(got):
        img, lab = src.batch(s)
        assert b["image"].dtype == torch.bfloat16 and b["image"].is_cuda
        assert torch.equal(b["image"], img) and torch.equal(b["label"], lab)
        # the trainer's own conversion is then a no-op
        assert ops.normalize_u8(src.host_batch(s)["image"].to(dev), torch.bfloat16).shape == b["image"].shape
