#!/usr/bin/env python
"""Which allocator gave this node its symmetric slab, and does it carry an NVSwitch multicast (NVLS) mapping?  Run under
torchrun; prints one JSON line from rank 0.  PICD_SYMM=own (default: the library's own cuMem* / cuMulticast* slab, then torch's
symmetric memory, then cudaMalloc + IPC) | torch | 0."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from picd_b200.parallel import dist as D  # noqa: E402
from picd_b200.parallel import native_comm  # noqa: E402

D.init_distributed("b200")
dev = torch.device("cuda", torch.cuda.current_device())
slab = native_comm.SymmetricSlab(1 << 20, 1, dev, dist.get_world_size(), dist.get_rank())
# a peer store followed by a read-back proves the peer mappings; a multimem.ld_reduce through the bucket kernel is exercised
# by benchmarks/multigpu_check.py (modes nvls / nvls_ld)
out = {"world": dist.get_world_size(), "backing": slab.backing, "multicast": bool(slab.mc_base),
       "nvls_supported": native_comm.nvls_supported(), "peer_bases": [hex(b) for b in slab.peer_bases]}
if dist.get_rank() == 0:
    print(json.dumps(out))
D.barrier()
D.shutdown()
