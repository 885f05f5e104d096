"""Is the W = 3 AdamW gap a kernel fault or AdamW's own sensitivity?  Two torch references that differ only in the ORDER the
ranks' gradients are added (0,1,2 vs 2,1,0: an fp32 last-bit effect) are compared with each other and with the kernel, per step."""
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
net, refA, refB = mk(), mk(), mk()
refA.load_state_dict(net.state_dict()); refB.load_state_dict(net.state_dict())
dp = FusedDataParallel(net, OptimizerSpec(kind="adamw", lr=1e-3, weight_decay=0.01), first_bucket_bytes=64 << 10,
                       bucket_cap_bytes=1 << 20, comm=os.environ.get("MODE", "two_shot"))
optA = torch.optim.AdamW(refA.parameters(), lr=1e-3, weight_decay=0.01)
optB = torch.optim.AdamW(refB.parameters(), lr=1e-3, weight_decay=0.01)


def reduce_into(model, order):
    for p in model.parameters():
        parts = [torch.empty_like(p.grad) for _ in range(world)]
        dist.all_gather(parts, p.grad.contiguous())
        tot = torch.zeros_like(p.grad)
        for r in order:
            tot += parts[r]
        p.grad.copy_(tot * (1.0 / world))


def gap(m1, m2):
    return max(float((a.detach() - b.detach()).abs().max() / b.detach().abs().max()) for a, b in zip(m1.parameters(), m2.parameters()))


for it in range(6):
    g = torch.Generator(device="cpu").manual_seed(100 * it + rank)
    x = torch.randn(32, 256, generator=g).to(device)
    dp(x).square().mean().backward()
    dp.optimizer.step()
    for ref, opt, order in ((refA, optA, list(range(world))), (refB, optB, list(range(world))[::-1])):
        opt.zero_grad()
        ref(x).square().mean().backward()
        reduce_into(ref, order)
        opt.step()
    torch.cuda.synchronize()
    if rank == 0:
        print(json.dumps({"world": world, "step": it + 1, "kernel_vs_refA": gap(net, refA), "refB_vs_refA": gap(refB, refA),
                          "kernel_vs_refB": gap(net, refB)}), flush=True)
dist.barrier()
