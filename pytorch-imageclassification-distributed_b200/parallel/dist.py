"""Process-group helpers: init / rank / world_size / barrier / reduce_tensor / all_gather.

API follows the reference's ``ddp_utils.py`` (``reduce_tensor`` :8-12, ``all_gather``
:16-56) plus the init/rank/barrier helpers BASELINE.json asks for.  Differences,
all fixes of SURVEY.md 0.1 defects:

* D2  ``--local_rank``, ``--local-rank`` and the ``LOCAL_RANK`` env var are all honoured
* D11 the global rank comes from ``RANK`` (multi-node safe), not from the local rank
* D14 nothing hard-codes ``"cuda"``: helpers follow the backend (nccl -> current CUDA
      device, gloo -> CPU) so the CPU plumbing config runs
"""
from __future__ import annotations

import datetime
import os
import pickle
from typing import Any, List, Optional

import torch
import torch.distributed as dist


def is_dist() -> bool:
    return dist.is_available() and dist.is_initialized()


def get_rank() -> int:
    return dist.get_rank() if is_dist() else 0


def get_world_size() -> int:
    return dist.get_world_size() if is_dist() else 1


def get_local_rank(cli_value: Optional[int] = None) -> int:
    """LOCAL_RANK env (torchrun) wins over the ``--local_rank`` flag (torch.distributed.launch)."""
    env = os.environ.get("LOCAL_RANK")
    if env is not None:
        return int(env)
    return int(cli_value or 0)


def is_main() -> bool:
    return get_rank() == 0


def comm_device() -> torch.device:
    """Device collectives must run on for the active backend."""
    if is_dist() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def init_distributed(backend: Optional[str] = None, local_rank: Optional[int] = None,
                     timeout_s: float = 600.0) -> int:
    """env:// rendezvous like train.py:102, one process per GPU.  Binds the CUDA device
    BEFORE creating the NCCL communicator (the reference binds after, train.py:106).
    Runs single-process (no group) when WORLD_SIZE is absent.  Returns the local rank."""
    lr = get_local_rank(local_rank)
    use_cuda = torch.cuda.is_available() and backend not in ("gloo",)
    if backend in (None, "b200"):
        backend = "nccl" if use_cuda else "gloo"
    if use_cuda:
        torch.cuda.set_device(lr % max(torch.cuda.device_count(), 1))
    if "WORLD_SIZE" in os.environ and not is_dist():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend, timeout=datetime.timedelta(seconds=timeout_s), **kw)
    return lr


def shutdown() -> None:
    if is_dist():
        dist.destroy_process_group()


def barrier() -> None:
    if is_dist():
        if dist.get_backend() == "nccl":
            dist.barrier(device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier()


def reduce_tensor(tensor: torch.Tensor, average: bool = True) -> torch.Tensor:
    """Sum (and by default average) a tensor over all ranks; the input is untouched."""
    rt = tensor.detach().clone()
    if get_world_size() == 1:
        return rt
    dev = comm_device()
    moved = rt.to(dev)
    dist.all_reduce(moved, op=dist.ReduceOp.SUM)
    if average:
        moved = moved / get_world_size() if not moved.is_floating_point() else moved.div_(get_world_size())
    return moved.to(tensor.device)


def all_gather(data: Any) -> List[Any]:
    """Gather one picklable object per rank, returned in rank order on every rank.

    Same contract as ddp_utils.py:16-56 (world size 1 short-circuits to ``[data]``); the
    transport is one size exchange + one padded byte exchange, on whatever device the
    backend needs.
    """
    world = get_world_size()
    if world == 1:
        return [data]
    dev = comm_device()
    payload = torch.frombuffer(bytearray(pickle.dumps(data)), dtype=torch.uint8).to(dev)
    sizes = torch.zeros(world, dtype=torch.int64, device=dev)
    sizes[get_rank()] = payload.numel()
    dist.all_reduce(sizes)                      # one collective instead of W item() syncs
    sizes_l = sizes.tolist()
    cap = max(sizes_l)
    buf = torch.zeros(cap, dtype=torch.uint8, device=dev)
    buf[: payload.numel()] = payload
    out = torch.empty(world * cap, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(out, buf)
    host = out.cpu().numpy()
    return [pickle.loads(host[r * cap: r * cap + sizes_l[r]].tobytes()) for r in range(world)]


def broadcast_object(obj: Any, src: int = 0) -> Any:
    if get_world_size() == 1:
        return obj
    box = [obj]
    dist.broadcast_object_list(box, src=src, device=comm_device())
    return box[0]
