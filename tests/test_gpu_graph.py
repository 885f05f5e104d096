"""CUDA-graph replay of the whole training step (engine/graphed.py): the replayed steps must train like the eagerly
launched ones, and the step-varying scalars (learning rate, Adam bias correction, dropout mask) must follow the host
values although the launches were recorded only once."""
import pytest

pytestmark = pytest.mark.gpu


def _run(graph: bool, steps: int = 10, kind: str = "sgd", name: str = "resnet18", lr_zero_from=None):
    import torch
    from picd_b200.engine.graphed import GraphedTrainStep
    from picd_b200.engine.trainer import WeightedCrossEntropy
    from picd_b200.models import Classifier
    from picd_b200.ops import reference as ref
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = Classifier(name, 7, image_size=64).to(dev)
    spec = OptimizerSpec(kind, lr=0.02 if kind == "sgd" else 1e-3, momentum=0.9, weight_decay=1e-4)
    dp = FusedDataParallel(model, spec)
    dp.train()
    loss_fn = WeightedCrossEntropy(torch.tensor([3, 3, 10, 1, 4, 4, 5.0], device=dev))
    step = GraphedTrainStep(dp, loss_fn, dp.optimizer, warmup=2, enabled=graph)
    losses, snaps = [], []
    for s in range(steps):
        if lr_zero_from is not None and s >= lr_zero_from:
            spec.lr = 0.0
        img, lab = ref.synthetic_batch(32, 3, 64, 64, 7, 0, s % 2, "cpu", torch.float32)
        losses.append(float(step(img.to(dev).to(torch.bfloat16), lab.to(dev)).item()))
        snaps.append(dp.arena.master[:4096].clone())
    return losses, snaps, step, dp


def test_graph_replay_trains_like_eager():
    eager, _, _, _ = _run(False)
    graphed, _, step, dp = _run(True)
    assert step.graph is not None and step.replays == 8 and step.launches_per_replay > 50
    assert dp.step_index == 10 and dp.optimizer.step_count == 10
    assert graphed[-1] < graphed[0]
    for a, b in zip(eager, graphed):
        assert abs(a - b) < 0.08 * max(1.0, abs(a)), (eager, graphed)


def test_graph_replay_reads_lr_from_device():
    """lr drops to 0 while the graph is being replayed: SGD momentum keeps moving the weights unless lr itself is
    re-read, so identical masters before/after a replay prove the kernel saw the new value."""
    import torch
    _, snaps, step, _ = _run(True, steps=8, lr_zero_from=5)
    assert step.replays == 6
    assert not torch.equal(snaps[3], snaps[4])          # lr > 0: weights move under replay
    assert torch.equal(snaps[5], snaps[6]) and torch.equal(snaps[6], snaps[7])


def test_graph_replay_adam_bias_correction():
    eager, _, _, _ = _run(False, kind="adam", steps=8)
    graphed, _, step, _ = _run(True, kind="adam", steps=8)
    assert step.replays == 6
    for a, b in zip(eager, graphed):
        assert abs(a - b) < 0.08 * max(1.0, abs(a)), (eager, graphed)


def test_graph_replay_dropout_mask_changes():
    """The recorded dropout launch mixes its seed with the device-side step counter: two replays draw different masks."""
    import torch
    from picd_b200 import ops
    from picd_b200.ops import native
    dev = torch.device("cuda", 0)
    word = torch.zeros(1, dtype=torch.float32, device=dev)
    x = torch.ones(1 << 14, dtype=torch.bfloat16, device=dev)
    native.set_step_word(word)
    try:
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            y = ops.dropout(x, 0.5, True)
        outs = []
        for i in range(3):
            word.view(torch.int32).fill_(i + 1)
            g.replay()
            outs.append(y.clone())
    finally:
        native.set_step_word(None)
    assert not torch.equal(outs[0], outs[1]) and not torch.equal(outs[1], outs[2])
    for o in outs:
        assert 0.4 < float((o == 0).float().mean()) < 0.6


def test_batch_counters_deferred_but_exact():
    """num_batches_tracked is bumped by one multi-tensor add per forward (eager and replayed) and is exact when read."""
    import torch
    _, _, step, dp = _run(True, steps=7)
    sd = dp.state_dict()
    counts = {k: int(v) for k, v in sd.items() if k.endswith("num_batches_tracked")}
    assert counts and set(counts.values()) == {7}, counts
    bn = dp.module.encoder.bn1
    dp.eval()
    with torch.no_grad():
        dp(torch.randn(2, 3, 64, 64, device="cuda").to(torch.bfloat16))
    assert int(bn.num_batches_tracked) == 7            # eval forwards do not count
