"""W = 3 AdamW, step 2: do the local gradients of the engine's weights and of the reference's weights differ, and if so where do the
weights differ?"""
import os, sys, json
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from picd_b200.parallel import init_distributed
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec

init_distributed("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
os.environ["PICD_GRAD_WIRE"] = "fp32"
torch.manual_seed(0)
ACT = getattr(torch.nn, os.environ.get("ACT", "ReLU"))
mk = lambda: torch.nn.Sequential(torch.nn.Linear(256, 512), ACT(), torch.nn.Linear(512, 1024), ACT(),
                                 torch.nn.Linear(1024, 8)).to(device)
net, ref, probe = mk(), mk(), mk()
ref.load_state_dict(net.state_dict())
dp = FusedDataParallel(net, OptimizerSpec(kind="adamw", lr=1e-3, weight_decay=0.01), first_bucket_bytes=64 << 10,
                       bucket_cap_bytes=1 << 20, comm="one_shot")
opt = torch.optim.AdamW(ref.parameters(), lr=1e-3, weight_decay=0.01)
for it in range(3):
    g = torch.Generator(device="cpu").manual_seed(100 * it + rank)
    x = torch.randn(32, 256, generator=g).to(device)
    probe.load_state_dict(net.state_dict())
    probe.zero_grad()
    probe(x).square().mean().backward()
    dp(x).square().mean().backward()
    dp.optimizer.step()
    opt.zero_grad()
    ref(x).square().mean().backward()
    rec = {"rank": rank, "step": it + 1}
    rec["local_grad_probe_vs_ref"] = [float((a.grad - b.grad).abs().max() / b.grad.abs().max()) for a, b in zip(probe.parameters(), ref.parameters())]
    rec["w_before_step_engine_vs_ref"] = [float((a.detach() - b.detach()).abs().max()) for a, b in zip(probe.parameters(), ref.parameters())]
    for p in ref.parameters():
        parts = [torch.empty_like(p.grad) for _ in range(world)]
        dist.all_gather(parts, p.grad.contiguous())
        tot = torch.zeros_like(p.grad)
        for part in parts:
            tot += part
        p.grad.copy_(tot * (1.0 / world))
    # same reduction of the probe's gradients = what the kernel should have seen
    red = []
    for p in probe.parameters():
        parts = [torch.empty_like(p.grad) for _ in range(world)]
        dist.all_gather(parts, p.grad.contiguous())
        tot = torch.zeros_like(p.grad)
        for part in parts:
            tot += part
        red.append(tot * (1.0 / world))
    rec["reduced_grad_probe_vs_ref"] = [float((a - b.grad).abs().max() / b.grad.abs().max()) for a, b in zip(red, ref.parameters())]
    opt.step()
    torch.cuda.synchronize()
    rec["w_after_step_engine_vs_ref"] = [float((a.detach() - b.detach()).abs().max()) for a, b in zip(net.parameters(), ref.parameters())]
    for k in list(rec):
        if isinstance(rec[k], list):
            rec[k] = float('%.2e' % max(rec[k]))
    print(json.dumps(rec), flush=True)
dist.barrier()
