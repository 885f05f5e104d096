"""Model zoo: state-dict parity with torchvision, forward parity on CPU, Classifier behaviour (nn/classifier.py:7-37)."""
import pytest
import torch
import torchvision

from picd_b200.models import (BACKBONES, Classifier, MLPHead, SyncBatchNorm, convert_sync_batchnorm, inception_v3,
                              resnet18, resnet50, vit_b_16)


def _same_keys(a, b):
    sa, sb = a.state_dict(), b.state_dict()
    assert set(sa) == set(sb)
    for k in sa:
        assert sa[k].shape == sb[k].shape, k


def test_state_dict_keys_match_torchvision():
    _same_keys(resnet18(), torchvision.models.resnet18())
    _same_keys(resnet50(), torchvision.models.resnet50())
    _same_keys(vit_b_16(), torchvision.models.vit_b_16())
    _same_keys(inception_v3(), torchvision.models.inception_v3(weights=None, aux_logits=True, init_weights=False))


@pytest.mark.parametrize("builder,tv,size", [(resnet18, torchvision.models.resnet18, 64),
                                             (resnet50, torchvision.models.resnet50, 64)])
def test_resnet_forward_matches_torchvision(builder, tv, size):
    torch.manual_seed(0)
    ref = tv()
    ours = builder()
    ours.load_state_dict(ref.state_dict())
    x = torch.randn(2, 3, size, size)
    for mode in ("train", "eval"):
        getattr(ref, mode)(); getattr(ours, mode)()
        assert torch.allclose(ours(x), ref(x), atol=2e-4, rtol=1e-4)
    # running statistics advanced identically in train mode
    assert torch.allclose(ours.bn1.running_var, ref.bn1.running_var, atol=1e-5)


def test_inception_forward_matches_torchvision_and_returns_aux():
    torch.manual_seed(0)
    ref = torchvision.models.inception_v3(weights=None, aux_logits=True, init_weights=False, transform_input=False)
    ours = inception_v3()
    ours.load_state_dict(ref.state_dict())
    x = torch.randn(2, 3, 299, 299)
    ref.eval(); ours.eval()
    assert torch.allclose(ours(x), ref(x), atol=1e-3, rtol=1e-3)
    ours.train()
    out = ours(x)
    assert isinstance(out, tuple) and out.logits.shape == (2, 1000) and out.aux_logits.shape == (2, 1000)


def test_vit_forward_matches_torchvision():
    torch.manual_seed(0)
    ref = torchvision.models.vit_b_16().eval()
    ours = vit_b_16().eval()
    ours.load_state_dict(ref.state_dict())
    x = torch.randn(1, 3, 224, 224)
    assert torch.allclose(ours(x), ref(x), atol=2e-3, rtol=1e-3)


def test_classifier_head_and_names():
    for name in ("resnet18", "resnet50", "inceptionv3", "efficientnet-b0", "vit_tiny_test"):
        m = Classifier(name, 7)
        head = m.encoder.heads.head if name.startswith("vit") else (m.encoder._fc if "efficientnet" in name else m.encoder.fc)
        assert isinstance(head, MLPHead)
        dims = [(l.in_features, l.out_features) for l in head if hasattr(l, "in_features")]
        assert [d[1] for d in dims] == [128, 64, 32, 7]
        assert [i for i, l in enumerate(head) if hasattr(l, "in_features")] == [0, 2, 4, 6]
    m = Classifier("inceptionv3", 7)
    assert m.encoder.AuxLogits.fc.out_features == 7 and m.returns_aux
    assert "encoder.fc.0.weight" in m.state_dict()


def test_classifier_rejects_unknown_and_empty():
    with pytest.raises(ValueError):
        Classifier("alexnet", 7)
    with pytest.raises(ValueError):
        Classifier("resnet18", 0)
    assert "resnet101" in BACKBONES and "efficientnet-b3" in BACKBONES


def test_convert_sync_batchnorm_marks_every_bn():
    m = convert_sync_batchnorm(Classifier("resnet18", 3, small_input=True))
    bns = [x for x in m.modules() if hasattr(x, "running_mean")]
    assert bns and all(isinstance(b, SyncBatchNorm) and b.sync_group is True for b in bns)
    # single process: falls back to local statistics and still trains
    out = m(torch.randn(2, 3, 32, 32))
    out.sum().backward()


def test_graphed_train_step_falls_back_to_eager_on_cpu():
    """GraphedTrainStep is the public step API of bench.py / train.py --cuda-graph; without CUDA it must simply run the
    eager sequence (and hand out detached tensors)."""
    import torch
    from picd_b200.engine.graphed import GraphedTrainStep
    from picd_b200.engine.trainer import WeightedCrossEntropy
    from picd_b200.models import Classifier
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    torch.manual_seed(0)
    dp = FusedDataParallel(Classifier("resnet18", 7, image_size=32, small_input=True), OptimizerSpec("sgd", lr=0.05, momentum=0.9))
    dp.train()
    step = GraphedTrainStep(dp, WeightedCrossEntropy(None), dp.optimizer, warmup=1)
    assert not step.enabled
    x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 7, (4,))
    before = dp.arena.master.clone()
    losses = [float(step(x, y)) for _ in range(3)]
    assert step.graph is None and step.replays == 0 and dp.step_index == 3
    assert not step(x, y).requires_grad and not step.outputs.requires_grad
    assert not torch.equal(before, dp.arena.master) and all(l == l for l in losses)


def test_cpu_fallbacks_of_small_ops():
    """concat_channels / top1_hits / fork_n run through the torch path on CPU tensors with the same semantics."""
    import torch
    from picd_b200 import ops
    from picd_b200.utils import Accuracy
    a = torch.randn(2, 8, 5, 5).contiguous(memory_format=torch.channels_last)
    b = torch.randn(2, 16, 5, 5).contiguous(memory_format=torch.channels_last)
    c = ops.concat_channels([a, b])
    assert torch.equal(c, torch.cat([a, b], 1)) and c.is_contiguous(memory_format=torch.channels_last)
    logits = torch.tensor([[0.1, 0.9, 0.0], [2.0, 1.0, 0.0], [0.0, 0.0, 0.5]])
    labels = torch.tensor([1, 1, 2])
    hits = ops.top1_hits(logits, labels)
    assert hits.tolist() == [1, 0, 1] and hits.dtype == torch.int64 and torch.equal(Accuracy(logits, labels), hits)
    x = torch.randn(1, 8, 4, 4, requires_grad=True)
    parts, fk = ops.fork_n(x, 3)
    assert fk is None and len(parts) == 3 and all(p is x for p in parts)
    x1, x2, fk2 = ops.fork(x)
    assert fk2 is None and x1 is x and x2 is x


def test_stem_op_cpu_equals_separate_ops():
    """ops.conv_bn_relu_maxpool (the ResNet stem as one op) on CPU == conv -> BN -> ReLU -> max_pool2d, values and gradients,
    and moves the running statistics the same way."""
    import torch
    import torch.nn.functional as F
    from picd_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(2, 3, 33, 33)
    w = torch.randn(16, 3, 7, 7) * 0.1
    g, b = torch.rand(16) + 0.5, torch.randn(16) * 0.1

    def run(fused):
        ps = [t.clone().requires_grad_() for t in (w, g, b)]
        rm, rv = torch.zeros(16), torch.ones(16)
        if fused:
            out = ops.conv_bn_relu_maxpool(x, ps[0], ps[1], ps[2], rm, rv, stride=2, padding=3, training=True)
        else:
            y = F.conv2d(x, ps[0], None, 2, 3)
            out = F.max_pool2d(F.relu(F.batch_norm(y, rm, rv, ps[1], ps[2], True, 0.1, 1e-5)), 3, 2, 1)
        out.square().sum().backward()
        return out.detach(), [p.grad for p in ps], rm, rv

    o1, g1, rm1, rv1 = run(True)
    o2, g2, rm2, rv2 = run(False)
    assert o1.shape == (2, 16, 9, 9)
    assert torch.allclose(o1, o2, atol=1e-5) and all(torch.allclose(a, c, atol=1e-4, rtol=1e-4) for a, c in zip(g1, g2))
    assert torch.allclose(rm1, rm2, atol=1e-6) and torch.allclose(rv1, rv2, atol=1e-6)


def test_channel_scale_cpu_fallback_and_stem_eval_mode():
    """ops.channel_scale on CPU == x * s[:, :, None, None] * mask (values and gradients); the stem op in eval mode uses the
    running statistics like the separate ops."""
    import torch
    import torch.nn.functional as F
    from picd_b200 import ops
    torch.manual_seed(1)
    x = torch.randn(3, 8, 5, 4, requires_grad=True)
    s = torch.rand(3, 8, requires_grad=True)
    m = torch.tensor([1.25, 0.0, 1.25])
    out = ops.channel_scale(x, s, m)
    out.sum().backward()
    xr, sr = x.detach().clone().requires_grad_(), s.detach().clone().requires_grad_()
    ref = xr * sr[:, :, None, None] * m.view(-1, 1, 1, 1)
    ref.sum().backward()
    assert torch.allclose(out, ref) and torch.allclose(x.grad, xr.grad) and torch.allclose(s.grad, sr.grad)
    assert torch.equal(ops.channel_scale(x, s), x * s[:, :, None, None])
    # stem, eval mode
    w = torch.randn(16, 3, 7, 7) * 0.1
    g, b = torch.rand(16) + 0.5, torch.randn(16) * 0.1
    rm, rv = torch.randn(16) * 0.1, torch.rand(16) + 0.5
    img = torch.randn(2, 3, 33, 33)
    with torch.no_grad():
        got = ops.conv_bn_relu_maxpool(img, w, g, b, rm.clone(), rv.clone(), stride=2, padding=3, training=False)
        want = F.max_pool2d(F.relu(F.batch_norm(F.conv2d(img, w, None, 2, 3), rm, rv, g, b, False, 0.1, 1e-5)), 3, 2, 1)
    assert torch.allclose(got, want, atol=1e-5)
