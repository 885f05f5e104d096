"""Worker run under torchrun (gloo, CPU) by tests/test_distributed_cpu.py: host-side data-parallel logic."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from picd_b200.engine.trainer import WeightedCrossEntropy, compute_loss  # noqa: E402
from picd_b200.models import Classifier, convert_sync_batchnorm  # noqa: E402
from picd_b200.ops import synthetic_batch  # noqa: E402
from picd_b200.parallel import all_gather, barrier, init_distributed, reduce_tensor, shutdown  # noqa: E402
from picd_b200.parallel.engine import FusedDataParallel, OptimizerSpec  # noqa: E402

init_distributed("gloo")
rank, world = dist.get_rank(), dist.get_world_size()

# ddp_utils parity: reduce_tensor averages, all_gather returns rank-ordered python objects (ddp_utils.py:8-56)
t = torch.tensor([float(rank + 1)])
assert reduce_tensor(t).item() == sum(range(1, world + 1)) / world and t.item() == rank + 1
objs = all_gather({"rank": rank, "arr": np.arange(rank + 2)})
assert [o["rank"] for o in objs] == list(range(world))
assert all(len(o["arr"]) == r + 2 for r, o in enumerate(objs))

for sync, kind in ((False, "sgd"), (True, "adam")):
    torch.manual_seed(1234 + rank)            # different init per rank: the engine must broadcast rank 0's weights
    model = Classifier("resnet18", 5, small_input=True)
    if sync:
        convert_sync_batchnorm(model)
    dp = FusedDataParallel(model, OptimizerSpec(kind, lr=0.05 if kind == "sgd" else 1e-3, momentum=0.9),
                           first_bucket_bytes=256 << 10, bucket_cap_bytes=4 << 20)
    assert len(dp.arena.buckets) > 2
    loss_fn = WeightedCrossEntropy(torch.tensor([1.0, 2.0, 1.0, 3.0, 1.0]))
    dp.train()
    losses = []
    for step in range(3):
        img, lab = synthetic_batch(4, 3, 32, 32, 5, 0, step * world + rank)       # each rank sees different data
        loss, _ = compute_loss(dp(img), lab, loss_fn)
        dp.optimizer.zero_grad()
        loss.backward()
        dp.optimizer.step()
        losses.append(loss.item())
    flat = dp.arena.master.clone()
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    assert all(torch.equal(gathered[0], g) for g in gathered), "replicas diverged"
    assert dp.arena.grad.abs().max().item() == 0.0, "gradient arena not re-zeroed"
    assert all(np.isfinite(l) for l in losses)
    if sync:   # SyncBN: running statistics identical on every rank without any buffer broadcast
        rv = dp.module.encoder.bn1.running_var.clone()
        rvs = [torch.empty_like(rv) for _ in range(world)]
        dist.all_gather(rvs, rv)
        assert all(torch.equal(rvs[0], r) for r in rvs)
# DDP's constructor check (torch/nn/parallel/distributed.py:881): replicas that disagree on the model raise on EVERY rank
# instead of corrupting each other's arenas
torch.manual_seed(0)
odd = Classifier("resnet18", 5 if rank == 0 else 6, small_input=True)
try:
    FusedDataParallel(odd, OptimizerSpec("sgd"))
    raise SystemExit("replica mismatch was not detected")
except RuntimeError as e:
    assert "replicas disagree" in str(e), e

# optimizer-state checkpoint is a collective that returns the full state on every rank, and reloads
sd = dp.optimizer.state_dict()
assert len(sd["state"]) == 2 and all(t.numel() == dp.arena.total for t in sd["state"])
dp.optimizer.load_state_dict(sd)

# descriptor exchange used by the VMM slab (native_comm.pass_fds): every rank ends up with descriptors of ITS OWN process
# that refer to the other ranks' kernel objects (memfd here; a cuMemCreate handle on the GPU box)
from picd_b200.parallel.native_comm import pass_fds  # noqa: E402

mine = os.memfd_create(f"picd_test_{rank}")
os.write(mine, f"rank{rank}".encode())
fds = pass_fds(rank, world, mine, range(world))
assert sorted(fds) == [r for r in range(world) if r != rank], fds
for r, fd in fds.items():
    assert os.pread(fd, 16, 0) == f"rank{r}".encode()
    os.close(fd)
one = pass_fds(rank, world, mine, [0])              # the multicast handle travels from rank 0 only
assert (rank == 0 and one == {}) or (rank != 0 and os.pread(one[0], 16, 0) == b"rank0")
os.close(mine)
barrier()
if rank == 0:
    print("DIST_WORKER_OK")
shutdown()
