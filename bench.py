#!/usr/bin/env python
"""Headline benchmark: ResNet-50 classifier training throughput (images/sec) at 224x224, bf16, SGD-momentum,
data-parallel over N GPUs of one node (BASELINE.json: "images/sec at 224^2 bf16 (device-timed, max over ranks)").

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \
        bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl reference ...        # the unmodified reference (baseline/_ref) on the same config

Protocol: W untimed warm-up steps, then exactly K steps bracketed by barrier + torch.cuda.synchronize() and a pair
of CUDA events on the compute stream; the reported time is the MAX over ranks.  Weak scaling: the per-GPU batch is
fixed, `value` is the whole-job images/sec.  Every step touches >1 GB of activations, i.e. inputs far larger than
the 126 MB L2, so no explicit L2 flush is needed between iterations.  SM clocks / throttle reasons are sampled with
nvidia-smi during the timed region.  `e2e` repeats the measurement through the public training API with the batch
crossing PCIe from pinned host memory every step (copied one step ahead on a copy stream, see data/prefetch.py) and the
loss read back to the host every step.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NUM_CLASSES = 7                           # the reference hard-codes 7 class weights (train.py:157)
CLASS_WEIGHTS = [3, 3, 10, 1, 4, 4, 5]
LR, MOMENTUM, WEIGHT_DECAY = 0.1, 0.9, 1e-4


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch")
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--sync-bn", action="store_true")
    ap.add_argument("--optimizer", default="sgd")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel from Python instead of replaying the "
                    "captured CUDA graph of the step")
    ap.add_argument("--local_rank", "--local-rank", type=int, default=0)
    return ap.parse_args()


def published_baseline():
    try:
        with open(os.path.join(ROOT, "BASELINE.json")) as f:
            pub = json.load(f).get("published", {})
        for v in pub.values():
            if isinstance(v, (int, float)):
                return float(v)
    except Exception:
        pass
    return None


def timed_region(run_steps, steps, device, rank):
    """barrier+sync, events around `steps` steps, barrier+sync; returns max-over-ranks milliseconds + clocks."""
    import torch
    import torch.distributed as dist
    from picd_b200.utils.clocks import ClockSampler
    multi = dist.is_available() and dist.is_initialized()
    if multi:
        dist.barrier(device_ids=[device.index])
    torch.cuda.synchronize()
    sampler = ClockSampler(device.index).start() if rank == 0 else None
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    run_steps(steps)
    b.record()
    torch.cuda.synchronize()
    if multi:
        dist.barrier(device_ids=[device.index])
    ms = torch.tensor([a.elapsed_time(b)], device=device)
    if multi:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    clocks = sampler.stop() if sampler is not None else None
    return float(ms.item()), clocks


# ------------------------------------------------------------------------------------------ our arm
def run_b200(args):
    import torch
    from picd_b200.data import DevicePrefetcher, SyntheticDeviceLoader, SyntheticHostLoader
    from picd_b200.engine.graphed import GraphedTrainStep
    from picd_b200.engine.trainer import WeightedCrossEntropy
    from picd_b200.models import Classifier, convert_sync_batchnorm
    from picd_b200.ops import require_lib
    from picd_b200.parallel import dist as D
    from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec

    local_rank = D.init_distributed("b200", args.local_rank)
    device = torch.device("cuda", torch.cuda.current_device())
    rank, world = D.get_rank(), D.get_world_size()
    lib = require_lib()
    lib.picd_launch_count.restype = __import__("ctypes").c_ulonglong
    torch.manual_seed(0)
    model = Classifier(args.model, NUM_CLASSES, image_size=args.image_size)
    if args.sync_bn:
        convert_sync_batchnorm(model)
    model = model.to(device)
    spec = OptimizerSpec(kind=args.optimizer, lr=LR if args.optimizer == "sgd" else 1e-3, momentum=MOMENTUM,
                         weight_decay=WEIGHT_DECAY)
    dp = FusedDataParallel(model, spec)
    dp.train()
    loss_fn = WeightedCrossEntropy(torch.tensor(CLASS_WEIGHTS, dtype=torch.float32, device=device))
    dev_loader = SyntheticDeviceLoader(args.batch, args.image_size, NUM_CLASSES, 1 << 30, device, seed=0)
    counter = {"i": 0}
    last = {}

    # the public training-step API: eager for the first calls, then one CUDA-graph replay per step (engine/graphed.py)
    step = GraphedTrainStep(dp, loss_fn, dp.optimizer, warmup=max(1, min(2, args.warmup - 1)), enabled=not args.no_graph)

    def run_device(n):
        for _ in range(n):
            images, labels = dev_loader.batch(counter["i"])      # generated + normalised on the device
            counter["i"] += 1
            last["loss"] = step(images, labels)

    run_device(args.warmup)
    l0, r0 = lib.picd_launch_count(), step.replays
    ms, clocks = timed_region(run_device, args.steps, device, rank)
    # kernels of this library launched in the timed region: direct launches + those inside the replayed graphs
    launches = int(lib.picd_launch_count() - l0) + (step.replays - r0) * step.launches_per_replay
    final_loss = float(last["loss"].item())
    value = args.batch * world * args.steps / (ms / 1e3)

    e2e = None
    if not args.no_e2e:
        host_loader = SyntheticHostLoader(args.batch, args.image_size, NUM_CLASSES, 1 << 30, device, seed=1)
        sink = {"loss": 0.0}
        from picd_b200.utils import LaggedScalarReader
        reader = LaggedScalarReader(device, lag=1)

        def run_e2e(n):
            # pinned uint8 batch -> H2D + device normalise on the copy stream (prefetched one step ahead, like the training
            # CLI does) -> training step -> D2H read of EVERY step's loss into pinned memory, consumed one step late so that
            # the host never waits for the step it has just enqueued (utils/lagged.py); the last value is read before
            # the region ends
            first = counter["i"]
            counter["i"] += n
            batches = (host_loader.host_batch(first + k) for k in range(n))
            for b in DevicePrefetcher(batches, device):
                v = reader.push(step(b["image"], b["label"]))
                if v is not None:
                    sink["loss"] = v
            for v in reader.flush():
                sink["loss"] = v

        run_e2e(max(4, args.warmup))          # every pinned buffer and the copy stream have been through a full cycle
        ms_e, _ = timed_region(run_e2e, args.steps, device, rank)
        e2e = {"value": args.batch * world * args.steps / (ms_e / 1e3), "unit": "images/sec",
               "h2d_bytes_per_step": host_loader.h2d_bytes_per_step, "d2h_bytes_per_step": 4,
               "ms_per_step": ms_e / args.steps}

    # replica identity (what DDP guarantees, train.py:128), checked OUTSIDE the timed regions: every rank hashes its
    # weights (BN running statistics are rank-local between two forwards unless SyncBN is on, exactly as under DDP's
    # per-forward buffer broadcast) and the hashes are compared across ranks
    replicas_identical = None
    if world > 1:
        import hashlib
        import torch.distributed as dist
        torch.cuda.synchronize()
        # the bf16 compute copies are what every rank trains with (the fp32 masters are sharded by slice owner)
        h = hashlib.sha1(dp.arena.shadow.detach().view(torch.int16).cpu().numpy().tobytes())
        if args.sync_bn:
            for b in dp.module.buffers():
                h.update(b.detach().cpu().numpy().tobytes())
        box = [None] * world
        dist.all_gather_object(box, h.hexdigest())
        replicas_identical = all(v == box[0] for v in box)

    if rank == 0:
        pub = published_baseline()
        print(json.dumps({
            "metric": "images/sec at 224^2 bf16 (device-timed, max over ranks)", "value": value, "unit": "images/sec",
            "impl": "b200", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": (value / pub) if pub else None,
            "dtype": "bf16", "data": "synthetic (generated + normalised on device; random-init weights)",
            "config": {"model": f"{args.model}+MLP head ({NUM_CLASSES} classes)", "global_batch": args.batch * world,
                       "per_gpu_batch": args.batch, "image_size": args.image_size, "seq_len": None,
                       "optimizer": (f"sgd-momentum lr={LR} wd={WEIGHT_DECAY}" if args.optimizer == "sgd"
                                     else f"{args.optimizer} lr=0.001 wd={WEIGHT_DECAY}"), "sync_bn": args.sync_bn,
                       "parallelism": f"dp{world}", "l2": "inputs larger than L2 (>1 GB activations per step)"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "final_loss": final_loss,
            "cuda_graph": bool(step.graph is not None), "replicas_identical": replicas_identical}))
    D.barrier()
    D.shutdown()


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """Runs the UNMODIFIED reference sources installed in baseline/_ref: its Classifier (nn/classifier.py), its
    train_epoch (train.py:36-73, compiled from the installed file without executing the module-level script), its
    AverageMeter, wrapped exactly as train.py:124-128 does (SyncBatchNorm optional, .cuda(), DDP).  Deviations, all
    forced: random-init weights instead of the download, synthetic batches, bf16 autocast and
    SGD-momentum as named by the BASELINE.json config, LOCAL_RANK handling."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "train.py")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is not installed (see DESIGN.md)"}))
        return
    try:
        from baseline.ref_bench import run
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"impl": "reference", "unavailable": f"harness import failed: {e!r}"[:200]}))
        return
    run(args, ref_dir, timed_region, published_baseline, dict(num_classes=NUM_CLASSES, class_weights=CLASS_WEIGHTS,
                                                              lr=LR, momentum=MOMENTUM, weight_decay=WEIGHT_DECAY))


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
