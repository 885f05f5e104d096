"""Reference arm of bench.py: the UNMODIFIED reference sources from ``baseline/_ref`` on the bench config.

What runs is the reference's own code: ``Classifier`` from ``nn/classifier.py``, ``AverageMeter`` from ``utils.py`` and
the ``train_epoch`` function of ``train.py`` (lines 36-73), compiled from the installed file by extracting that one
function definition -- importing ``train.py`` is impossible because its module level parses argv, initialises NCCL
and starts a 100-epoch job (train.py:27-31,99-188).  The wrapping follows train.py:124-128: optional
``SyncBatchNorm.convert_sync_batchnorm``, ``.cuda()``, optimizer on the bare module's parameters, then
``DistributedDataParallel(device_ids=[local_rank], output_device=local_rank, find_unused_parameters=False)``.

Forced deviations (SURVEY.md 0.1 / BASELINE.md): ``pretrained=True`` cannot download, so the torchvision builders
are asked for random-init weights; ``efficientnet_pytorch`` is not installed, so an empty stand-in module satisfies
the unconditional import; batches are synthetic; bf16 autocast + SGD-momentum are the BASELINE.json config (the
reference itself is fp32 + Adam); LOCAL_RANK comes from the environment.  None of picd_b200's models, kernels or
engine are on this path.
"""
from __future__ import annotations

import argparse
import ast
import importlib.util
import json
import os
import socket
import sys
import types


def _load(path: str, name: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run(args, ref_dir, timed_region, published_baseline, cfg):
    import torch
    import torch.distributed as dist
    import torch.nn as tnn
    import torchvision.models as tvm
    from torch.nn.parallel import DistributedDataParallel as DDP
    from tqdm import tqdm

    local_rank = int(os.environ.get("LOCAL_RANK", args.local_rank))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if "WORLD_SIZE" not in os.environ:                 # plain `python bench.py --impl reference --gpus 1`
        os.environ.update(RANK="0", WORLD_SIZE="1", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=device)   # train.py:102
    rank, world = dist.get_rank(), dist.get_world_size()

    # environment shims (the reference files themselves are untouched)
    stub = types.ModuleType("efficientnet_pytorch")
    stub.EfficientNet = type("EfficientNet", (), {})
    sys.modules.setdefault("efficientnet_pytorch", stub)
    for name in ("resnet50", "resnet101", "inception_v3"):
        orig = getattr(tvm, name)
        setattr(tvm, name, (lambda o: (lambda pretrained=False, **kw: o(weights=None, **kw)))(orig))

    ref_classifier = _load(os.path.join(ref_dir, "nn", "classifier.py"), "ref_nn_classifier")
    ref_utils = _load(os.path.join(ref_dir, "utils.py"), "ref_utils")
    train_py = os.path.join(ref_dir, "train.py")
    tree = ast.parse(open(train_py).read())
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "train_epoch"][0]
    ns = {"torch": torch, "dist": dist, "tqdm": tqdm, "AverageMeter": ref_utils.AverageMeter,
          "args": argparse.Namespace(local_rank=local_rank)}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), train_py, "exec"), ns)
    train_epoch = ns["train_epoch"]

    name = {"resnet50": "resnet50", "resnet101": "resnet101", "inceptionv3": "inceptionv3"}.get(args.model)
    note = None
    torch.manual_seed(0)
    if name is not None:
        model = ref_classifier.Classifier(name=name, num_classes=cfg["num_classes"])
    elif args.model == "vit_b_16":
        # BASELINE.json names ViT-B/16 as a target, but the reference's Classifier has no ViT branch
        # (nn/classifier.py:11-23).  The closest reference-style arm: the reference's train_epoch + DDP wrapping around a
        # torchvision ViT-B/16 whose head is replaced by the reference's MLP head, attribute name `encoder` as in Classifier.
        name = "vit_b_16"
        note = "reference train_epoch + DDP around torchvision vit_b_16 with the reference's MLP head (its Classifier has no ViT)"

        class _RefStyleViT(tnn.Module):
            def __init__(self, nc):
                super().__init__()
                self.encoder = tvm.vit_b_16(weights=None, image_size=args.image_size)
                nf = self.encoder.heads.head.in_features
                self.encoder.heads = tnn.Sequential(tnn.Linear(nf, 128), tnn.ReLU(inplace=True), tnn.Linear(128, 64),
                                                    tnn.ReLU(inplace=True), tnn.Linear(64, 32), tnn.ReLU(inplace=True),
                                                    tnn.Linear(32, nc))

            def forward(self, images):
                return self.encoder(images)

        model = _RefStyleViT(cfg["num_classes"])
    else:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"the reference has no backbone named {args.model}"}))
        dist.destroy_process_group()
        return
    if args.sync_bn:
        model = tnn.SyncBatchNorm.convert_sync_batchnorm(model, None)              # train.py:124
    model = model.cuda()
    if args.optimizer == "adamw":
        optimizer = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=cfg["weight_decay"])
    elif args.optimizer == "adam":
        optimizer = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=cfg["weight_decay"])
    else:
        optimizer = torch.optim.SGD(model.parameters(), lr=cfg["lr"], momentum=cfg["momentum"],
                                    weight_decay=cfg["weight_decay"])
    model = DDP(model, device_ids=[local_rank], output_device=local_rank, find_unused_parameters=False)  # train.py:128
    weight = torch.tensor(cfg["class_weights"], requires_grad=False, dtype=torch.float32).cuda()
    loss_fn = tnn.CrossEntropyLoss(weight=weight)                                   # train.py:157-158

    B, S = args.batch, args.image_size
    g = torch.Generator().manual_seed(1234 + rank)

    def make(pinned: bool, n: int):
        out = []
        for _ in range(n):
            img = torch.randn(B, 3, S, S, generator=g)
            lab = torch.randint(0, cfg["num_classes"], (B,), generator=g)
            if pinned:
                out.append({"image": img.pin_memory(), "label": lab.pin_memory()})
            else:
                out.append({"image": img.to(device), "label": lab.to(device)})
        return out

    def epoch(batches, n):
        seq = [batches[i % len(batches)] for i in range(n)]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            train_epoch(0, model, loss_fn, optimizer, seq, name)

    dev_batches = make(False, 2)
    epoch(dev_batches, args.warmup)
    ms, clocks = timed_region(lambda n: epoch(dev_batches, n), args.steps, device, rank)
    value = B * world * args.steps / (ms / 1e3)
    e2e = None
    if not args.no_e2e:
        host_batches = make(True, 3)
        epoch(host_batches, max(3, args.warmup // 2))
        ms_e, _ = timed_region(lambda n: epoch(host_batches, n), args.steps, device, rank)
        e2e = {"value": B * world * args.steps / (ms_e / 1e3), "unit": "images/sec",
               "h2d_bytes_per_step": B * 3 * S * S * 4 + B * 8, "d2h_bytes_per_step": 8,
               "ms_per_step": ms_e / args.steps}
    if rank == 0:
        pub = published_baseline()
        print(json.dumps({
            "metric": "images/sec at 224^2 bf16 (device-timed, max over ranks)", "value": value, "unit": "images/sec",
            "impl": "reference", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / pub) if pub else None, "dtype": "bf16 (torch.autocast)",
            "data": "synthetic (random tensors; random-init weights)",
            "config": {"model": f"{args.model}+MLP head ({cfg['num_classes']} classes)", "global_batch": B * world,
                       "per_gpu_batch": B, "image_size": S, "seq_len": None,
                       "optimizer": (f"sgd-momentum lr={cfg['lr']} wd={cfg['weight_decay']}" if args.optimizer == "sgd"
                                     else f"{args.optimizer} lr=0.001 wd={cfg['weight_decay']}"), "sync_bn": args.sync_bn,
                       "parallelism": f"dp{world}", "l2": "inputs larger than L2 (>1 GB activations per step)"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": None, "note": note}))
    dist.barrier(device_ids=[local_rank])
    dist.destroy_process_group()
