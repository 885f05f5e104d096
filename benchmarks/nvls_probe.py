#!/usr/bin/env python
"""Does torch's symmetric-memory allocator give this node an NVSwitch multicast (NVLS) mapping?  Run under torchrun."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rank = int(os.environ.get("RANK", 0))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
out = {"rank": rank, "world": dist.get_world_size()}
try:
    import torch.distributed._symmetric_memory as sm
    t = sm.empty(1 << 22, dtype=torch.uint8, device=dev)
    h = sm.rendezvous(t, dist.group.WORLD)
    out.update(buffer_ptrs=[hex(p) for p in h.buffer_ptrs], multicast_ptr=hex(h.multicast_ptr), backend=str(sm.get_backend(dev)),
               signal_pad_size=h.signal_pad_size, local_ptr=hex(t.data_ptr()))
except Exception as e:  # noqa: BLE001
    out["error"] = repr(e)[:400]
print(json.dumps(out), flush=True)
dist.barrier(device_ids=[local])
dist.destroy_process_group()
