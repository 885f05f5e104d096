"""``train.py`` command line (reference train.py:27-31,99-188).

Kept: ``--local_rank`` / ``--datadir`` / ``--batchsize`` and the launch recipe
``python -m torch.distributed.launch --nproc_per_node=N train.py --datadir D`` (README.md:6).
Added (BASELINE.json north-star + SURVEY.md 7.1): ``--local-rank`` alias and ``LOCAL_RANK``
env (fix D2), ``--lr --epochs --model --optimizer --dtype --image-size --synthetic --sync-bn
--backend --ckpt-dir --class-weights``.  Every constant the reference hard-codes (image size
299, inceptionv3, Adam lr 0.5e-5, MultiStepLR [50, 80] x0.5, 100 epochs, class weights) is the
default of a flag, so a bare invocation trains the same recipe.
"""
from __future__ import annotations

import argparse
import os
from typing import List, Optional

import torch

from ..data import ImageDataset, SyntheticDeviceLoader, make_loader
from ..models import BACKBONES, Classifier, convert_sync_batchnorm
from ..parallel import ShardedSampler
from ..parallel import dist as D
from ..parallel.engine import FusedDataParallel, MultiStepLR, OptimizerSpec
from ..utils.checkpoint import BEST, load_checkpoint
from ..utils.tracing import Tracer
from .trainer import WeightedCrossEntropy, fit

REFERENCE_CLASS_WEIGHTS = [3, 3, 10, 1, 4, 4, 5]          # train.py:157


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="B200-native distributed image-classification trainer")
    p.add_argument("--local_rank", "--local-rank", dest="local_rank", default=0, type=int)
    p.add_argument("--datadir", default="", help="<datadir>/{train,valid}/<class>/*; optional with --synthetic")
    p.add_argument("--batchsize", "--batch-size", dest="batchsize", default=4, type=int)
    p.add_argument("--lr", default=0.5e-5, type=float)
    p.add_argument("--epochs", default=100, type=int)
    p.add_argument("--model", default="inceptionv3", choices=BACKBONES)
    p.add_argument("--optimizer", default="adam", choices=["sgd", "adam", "adamw"])
    p.add_argument("--momentum", default=0.9, type=float)
    p.add_argument("--weight-decay", default=0.0, type=float)
    p.add_argument("--milestones", default="50,80")
    p.add_argument("--gamma", default=0.5, type=float)
    p.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    p.add_argument("--image-size", default=None, type=int, help="default 299 (inception) / 224")
    p.add_argument("--synthetic", action="store_true", help="on-device synthetic batches")
    p.add_argument("--synthetic-steps", default=100, type=int)
    p.add_argument("--num-classes", default=None, type=int)
    p.add_argument("--class-weights", default=None, help="comma list; default = reference's 7 weights "
                   "when the dataset has 7 classes, else unweighted (fix D13)")
    p.add_argument("--sync-bn", action="store_true")
    p.add_argument("--backend", default="b200", choices=["b200", "nccl", "gloo"])
    p.add_argument("--ckpt-dir", default=os.path.join("dtmodel", "cp"))
    p.add_argument("--num-workers", default=6, type=int)
    p.add_argument("--log-interval", default=50, type=int)
    p.add_argument("--max-steps", default=None, type=int, help="cap steps per epoch (smoke runs)")
    p.add_argument("--pretrained", default=None, help="local torchvision state-dict file")
    p.add_argument("--trace", default=None, help="write phase timings / metrics JSONL here")
    p.add_argument("--seed", default=0, type=int)
    p.add_argument("--calibrate-comm", action="store_true",
                   help="measure one-shot vs two-shot all-reduce at start-up and pick the switch-over bucket size")
    p.add_argument("--cuda-graph", dest="cuda_graph", action="store_true", default=None,
                   help="record the training step once and replay it as a CUDA graph (default on the b200 backend on "
                        "CUDA, SyncBatchNorm included; batches of another shape, e.g. a ragged last one, run eagerly)")
    p.add_argument("--no-cuda-graph", dest="cuda_graph", action="store_false")
    p.add_argument("--progress", action="store_true",
                   help="rank 0 shows the reference's tqdm bar with the running loss instead of printed lines")
    return p


def _class_weights(args, num_classes: int, device) -> Optional[torch.Tensor]:
    if args.class_weights:
        w = [float(v) for v in args.class_weights.split(",")]
        if len(w) != num_classes:
            raise SystemExit(f"--class-weights has {len(w)} entries, dataset has {num_classes} classes")
        return torch.tensor(w, dtype=torch.float32, device=device)
    if num_classes == len(REFERENCE_CLASS_WEIGHTS):
        return torch.tensor(REFERENCE_CLASS_WEIGHTS, dtype=torch.float32, device=device)
    return None


def main(argv: Optional[List[str]] = None) -> int:
    args = build_parser().parse_args(argv)
    local_rank = D.init_distributed(args.backend, args.local_rank)
    use_cuda = torch.cuda.is_available() and args.backend != "gloo"
    device = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
    if args.dtype == "fp32" and use_cuda:
        # the sm_100a kernels are bf16-in / fp32-accumulate (fp32 masters, statistics and optimizer); there is no
        # fp32/tf32 tensor-core path, and silently training in bf16 under an fp32 flag would be a lie
        raise SystemExit("--dtype fp32 is only available on the CPU/gloo path (--backend gloo); the CUDA kernels "
                         "compute in bf16 with fp32 accumulation, masters and optimizer state")
    if args.cuda_graph is None:
        args.cuda_graph = use_cuda and args.backend == "b200"
    torch.manual_seed(args.seed)
    image_size = args.image_size or (299 if args.model == "inceptionv3" else 224)
    small = image_size <= 64

    if args.synthetic:
        num_classes = args.num_classes or 7
        train_loader = SyntheticDeviceLoader(args.batchsize, image_size, num_classes, args.synthetic_steps,
                                             device, seed=args.seed)
        val_loader = SyntheticDeviceLoader(args.batchsize, image_size, num_classes,
                                           max(1, args.synthetic_steps // 10), device, seed=args.seed + 1)
    else:
        if not args.datadir:
            raise SystemExit("--datadir is required unless --synthetic is given")
        raw = device.type == "cuda"
        train_ds = ImageDataset(args.datadir, "train", image_size, seed=args.seed, raw_u8=raw)
        val_ds = ImageDataset(args.datadir, "valid", image_size, seed=args.seed, raw_u8=raw,
                              class_names=sorted(train_ds.mapping))
        num_classes = args.num_classes or train_ds.num_classes
        train_loader = make_loader(train_ds, args.batchsize, ShardedSampler(train_ds), args.num_workers)
        val_loader = make_loader(val_ds, 1 if args.batchsize == 4 else args.batchsize,
                                 ShardedSampler(val_ds), args.num_workers)      # train.py:118: batch_size=1
        if device.type == "cuda":      # H2D copy + uint8 normalisation of batch i+1 overlap step i (data/prefetch.py)
            from ..data import DevicePrefetcher
            train_loader, val_loader = DevicePrefetcher(train_loader, device), DevicePrefetcher(val_loader, device)

    model = Classifier(args.model, num_classes, pretrained=args.pretrained, image_size=image_size,
                       small_input=small)
    if args.sync_bn:
        convert_sync_batchnorm(model)                                            # train.py:124
    model = model.to(device)
    spec = OptimizerSpec(kind=args.optimizer, lr=args.lr, momentum=args.momentum,
                         weight_decay=args.weight_decay)
    if args.calibrate_comm and device.type == "cuda" and D.get_world_size() > 1:
        from ..parallel import native_comm
        thr = native_comm.calibrate_two_shot_threshold(device)
        if D.is_main():
            print(f"comm: two-shot all-reduce from {thr} bytes (NVLS supported: {native_comm.nvls_supported()})")
    model = FusedDataParallel(model, spec)                                       # train.py:126-128
    optimizer = model.optimizer

    start_epoch, best_score, blob = load_checkpoint(model, args.ckpt_dir, args.model, BEST)   # train.py:131-153
    if blob:
        model.arena.refresh_shadow()
        if D.is_main():
            print(f"Loaded Checkpoint: {BEST}, with epoch {start_epoch} and best score {best_score}")
        start_epoch += 1                                                         # resume AFTER the saved epoch (fix D10)
    scheduler = MultiStepLR(optimizer, [int(m) for m in args.milestones.split(",") if m], args.gamma,
                            last_epoch=start_epoch)
    loss_fn = WeightedCrossEntropy(_class_weights(args, num_classes, device))
    tracer = Tracer(enabled=bool(args.trace), timing=True, jsonl_path=args.trace)

    history, best = fit(model, loss_fn, optimizer, scheduler, train_loader, val_loader, device,
                        name=args.model, epochs=args.epochs, start_epoch=start_epoch, best_score=best_score,
                        ckpt_dir=args.ckpt_dir, log_interval=args.log_interval, tracer=tracer,
                        max_steps=args.max_steps, cuda_graph=args.cuda_graph, progress=args.progress)
    if args.trace:
        tracer.log(event="summary", phases=tracer.summary(), history=history, best=best)
        tracer.close()
    D.barrier()
    D.shutdown()
    return 0
