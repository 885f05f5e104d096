"""The bucket kernel on ONE GPU (mode 0: local SGD-momentum / AdamW update + bf16 shadow + gradient re-zero) over a 256 MiB
bucket: what ncu can see of fused_allreduce_update_kernel (ncu is single-process; the peer phases are timed in sweep_w*.jsonl)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec

dev = torch.device("cuda")
for kind in ("sgd", "adamw"):
    net = torch.nn.Linear(8192, 8192, bias=False).to(dev)          # 64 Mi parameters = 256 MiB of fp32
    dp = FusedDataParallel(net, OptimizerSpec(kind=kind, lr=1e-3, momentum=0.9, weight_decay=0.01), first_bucket_bytes=1 << 30,
                           bucket_cap_bytes=1 << 30)
    for _ in range(2):
        dp.arena.grad.normal_()
        dp.optimizer.step()
    torch.cuda.synchronize()
    del dp, net
print("ok")
