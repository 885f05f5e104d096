#!/usr/bin/env python
"""Training-step time of every backbone on the native path (1 GPU, bf16, synthetic data): eagerly launched and replayed
from the captured CUDA graph of the step (engine/graphed.py)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from picd_b200.engine.graphed import GraphedTrainStep  # noqa: E402
from picd_b200.engine.trainer import WeightedCrossEntropy  # noqa: E402
from picd_b200.models import Classifier  # noqa: E402
from picd_b200.ops import synthetic_batch  # noqa: E402
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec  # noqa: E402

dev = torch.device("cuda")
out = []
ONLY = set(sys.argv[1:])        # optional: model names to run
for name, batch, size, opt in [("resnet18", 256, 224, "sgd"), ("resnet50", 256, 224, "sgd"), ("resnet101", 128, 224, "sgd"),
                               ("inceptionv3", 128, 299, "adam"), ("efficientnet-b3", 64, 300, "sgd"), ("vit_b_16", 128, 224, "adamw")]:
    if ONLY and name not in ONLY:
        continue
    try:
        torch.manual_seed(0)
        m = Classifier(name, 7, image_size=size).to(dev)
        dp = FusedDataParallel(m, OptimizerSpec(opt, lr=1e-3, momentum=0.9))
        dp.train()
        img, lab = synthetic_batch(batch, 3, size, size, 7, 0, 0, dev, torch.bfloat16)
        loss_fn = WeightedCrossEntropy(None)

        rec_ms = {}
        for mode in ("eager", "graph"):
            runner = GraphedTrainStep(dp, loss_fn, dp.optimizer, warmup=2, enabled=(mode == "graph"))
            for _ in range(4):
                l = runner(img, lab)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(8):
                l = runner(img, lab)
            b.record()
            torch.cuda.synchronize()
            rec_ms[mode] = a.elapsed_time(b) / 8
            if mode == "graph" and runner.graph is None:
                rec_ms[mode] = None
        ms = rec_ms["graph"] or rec_ms["eager"]
        rec = {"model": name, "batch": batch, "size": size, "optimizer": opt, "ms_per_step": ms, "ms_eager": rec_ms["eager"], "ms_graph": rec_ms["graph"], "img_per_s": batch / ms * 1e3,
               "loss": float(l.item()), "mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}
    except Exception as e:  # noqa: BLE001
        rec = {"model": name, "error": repr(e)[:300]}
    out.append(rec)
    print(json.dumps(rec), flush=True)
    del m, dp
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
open(os.path.join(ROOT, "gpurun_out", "model_steps.jsonl"), "w").write("\n".join(json.dumps(r) for r in out) + "\n")
