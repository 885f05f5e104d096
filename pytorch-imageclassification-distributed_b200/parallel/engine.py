"""The data-parallel engine: flat parameter arena, reverse-order gradient buckets and the
fused all-reduce + optimizer step.

What the reference does (train.py:126-128,71-73): ``Adam`` over the module's parameters,
``DistributedDataParallel`` (C++ Reducer, 1 MiB first bucket then 25 MiB buckets, NCCL
all-reduce from autograd hooks, average by 1/W), then a separate ``optimizer.step()`` that
re-reads every gradient and parameter from HBM.

What this engine does instead:

* all trainable parameters live in ONE flat fp32 *master* buffer, their gradients in ONE
  flat fp32 *gradient arena*, their bf16 compute copies in ONE flat *shadow* buffer
  (``param.data`` / ``param.grad`` / ``param._shadow`` are views).  The arena is laid out in
  reverse registration order, so every bucket is a contiguous range.
* on CUDA the arena is a ``cudaMalloc`` slab exported to the peer ranks with CUDA IPC; one
  kernel per bucket (launched on a side stream as soon as the bucket's last gradient is
  written) loads all W peers' gradient slices over NVLink, sums them in rank order, applies
  1/W, runs SGD-momentum / Adam / AdamW on the fp32 master, writes the bf16 shadow, pushes
  the result to the peers (two-shot) or keeps it local (one-shot), and re-zeroes the
  gradients -- there is no separate optimizer pass and no NCCL call on the hot path.
* on CPU/gloo (BASELINE.json config 1) the same bucket plan drives
  ``dist.all_reduce(async_op=True)`` and the torch oracle of the same update rule.

``FusedDataParallel`` keeps the wrapped network in ``.module`` so state-dict keys carry the
``module.`` prefix of the reference's checkpoints (train.py:176-180).
"""
from __future__ import annotations

import contextlib
import math
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import reference as ref
from .dist import get_rank, get_world_size, is_dist

ALIGN = 64                     # elements; keeps every view 256 B (fp32) / 128 B (bf16) aligned for TMA
FIRST_BUCKET_BYTES = 1 << 20   # torch DDP: dist._DEFAULT_FIRST_BUCKET_BYTES
BUCKET_CAP_BYTES = 25 << 20    # torch DDP default bucket_cap_mb=25


# ----------------------------------------------------------------------------- bucket plan
@dataclass
class Bucket:
    index: int
    start: int                 # element offset into the flat arena
    end: int
    params: List[int] = field(default_factory=list)     # indices into arena.params
    pending: int = 0
    launched: bool = False
    work: object = None

    @property
    def numel(self) -> int:
        return self.end - self.start


def plan_buckets(padded_sizes: Sequence[int], first_bytes: int = FIRST_BUCKET_BYTES,
                 cap_bytes: int = BUCKET_CAP_BYTES, elem_bytes: int = 4, unit: int = 1,
                 native: Optional[bool] = None) -> List[Tuple[int, int, List[int]]]:
    """Greedy contiguous bucketing over parameters already sorted in *expected gradient
    arrival order* (reverse registration).  The first bucket closes at ``first_bytes`` so
    communication starts early, later ones at ``cap_bytes``.  Every bucket's
    end is rounded up to a multiple of ``unit`` elements (``lcm(64, 4 * world)``): the
    two-shot kernel splits a bucket into ``world`` slices of whole 16-byte vectors, so a
    bucket must divide evenly for ANY world size (3, 5, 6, 7 included), not only powers of
    two.  Returns ``(start_elem, end_elem, [param indices])`` per bucket; the gap between a
    bucket's last parameter and its end is padding.

    ``csrc/runtime/bucket_planner.cpp`` (``picd_plan_buckets``) is the native implementation
    and is the one used whenever the kernel library is loadable; this Python body is the
    fallback for a tree without a build and the oracle of ``tests/test_bucket_planner.py``."""
    if native is None or native:
        out = _plan_buckets_native(padded_sizes, first_bytes, cap_bytes, elem_bytes, unit)
        if out is not None:
            return out
        if native:
            raise RuntimeError("picd_b200: native bucket planner requested but the library is not loadable")
    out: List[Tuple[int, int, List[int]]] = []
    start, cur, members = 0, 0, []
    limit = first_bytes
    for i, n in enumerate(padded_sizes):
        members.append(i)
        cur += n
        if (cur - start) * elem_bytes >= limit:
            cur = start + _round_up(cur - start, unit)
            out.append((start, cur, members))
            start, members, limit = cur, [], cap_bytes
    if members:
        cur = start + _round_up(cur - start, unit)
        out.append((start, cur, members))
    return out


def _plan_buckets_native(padded_sizes, first_bytes, cap_bytes, elem_bytes, unit):
    """``picd_plan_buckets`` through ctypes; None when the library cannot be loaded (CPU-only checkout without a build)."""
    import ctypes
    from ..ops._lib import lib_available, require_lib
    if not lib_available():
        return None
    lib = require_lib()
    if not hasattr(lib, "picd_plan_buckets"):
        return None
    n = len(padded_sizes)
    if n == 0:
        return []
    f = lib.picd_plan_buckets
    i64p = ctypes.POINTER(ctypes.c_int64)
    f.argtypes = [i64p, ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_int64, i64p, i64p,
                  ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    f.restype = ctypes.c_int
    sizes = (ctypes.c_int64 * n)(*[int(v) for v in padded_sizes])
    starts, ends = (ctypes.c_int64 * n)(), (ctypes.c_int64 * n)()
    owner = (ctypes.c_int * n)()
    nb = f(sizes, n, int(first_bytes), int(cap_bytes), int(elem_bytes), int(max(unit, 1)), starts, ends, owner, n)
    if nb < 0:
        raise RuntimeError("picd_plan_buckets failed")
    members: List[List[int]] = [[] for _ in range(nb)]
    for i in range(n):
        members[owner[i]].append(i)
    return [(int(starts[b]), int(ends[b]), members[b]) for b in range(nb)]


def _round_up(n: int, a: int) -> int:
    return (n + a - 1) // a * a


# ----------------------------------------------------------------------------- optimizer spec
@dataclass
class OptimizerSpec:
    """Update rule + hyper-parameters.  ``kind``: 'sgd' (momentum), 'adam' (reference default,
    train.py:127 ``Adam(lr=0.5e-5)``) or 'adamw'."""
    kind: str = "adam"
    lr: float = 0.5e-5
    momentum: float = 0.9
    nesterov: bool = False
    betas: Tuple[float, float] = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 0.0

    KINDS = {"sgd": 0, "adam": 1, "adamw": 2}

    def __post_init__(self):
        if self.kind not in self.KINDS:
            raise ValueError(f"optimizer kind {self.kind!r} not in {sorted(self.KINDS)}")

    @property
    def n_state(self) -> int:
        return 1 if self.kind == "sgd" else 2


class MultiStepLR:
    """``lr = base * gamma ** (#milestones <= epoch)``, stepped once per epoch
    (train.py:156,166: milestones [50, 80], gamma 0.5)."""

    def __init__(self, optimizer, milestones=(50, 80), gamma=0.5, last_epoch: int = 0):
        self.optimizer, self.milestones, self.gamma = optimizer, sorted(milestones), gamma
        self.base_lr, self.last_epoch = optimizer.spec.lr, last_epoch
        self._apply()

    def _apply(self):
        k = sum(1 for m in self.milestones if m <= self.last_epoch)
        self.optimizer.spec.lr = self.base_lr * self.gamma ** k

    def step(self):
        self.last_epoch += 1
        self._apply()

    def get_last_lr(self):
        return [self.optimizer.spec.lr]

    def state_dict(self):
        return {"last_epoch": self.last_epoch, "base_lr": self.base_lr}

    def load_state_dict(self, sd):
        self.last_epoch, self.base_lr = sd["last_epoch"], sd["base_lr"]
        self._apply()


# ----------------------------------------------------------------------------- arena
class ParamArena:
    """Flat master / gradient / shadow / optimizer-state buffers with per-parameter views."""

    def __init__(self, params: Iterable[nn.Parameter], device: torch.device, spec: OptimizerSpec,
                 first_bytes=FIRST_BUCKET_BYTES, cap_bytes=BUCKET_CAP_BYTES, world: int = 1, rank: int = 0,
                 buffer_floats: int = 0):
        plist = [p for p in params if p.requires_grad]
        self.params: List[nn.Parameter] = list(reversed(plist))     # expected gradient arrival order
        self.device, self.spec, self.world, self.rank = device, spec, world, rank
        sizes = [_round_up(p.numel(), ALIGN) for p in self.params]
        # every bucket must split into `world` slices of whole 16-byte vectors (two-shot / NVLS kernels), for any world size
        self.unit = math.lcm(ALIGN, 4 * max(world, 1))
        plan = plan_buckets(sizes, first_bytes, cap_bytes, unit=self.unit)
        self.buckets = [Bucket(i, s, e, m) for i, (s, e, m) in enumerate(plan)]
        self.offsets = [0] * len(self.params)
        for b in self.buckets:
            off = b.start
            for pi in b.params:
                self.offsets[pi] = off
                off += sizes[pi]
        self.total = self.buckets[-1].end if self.buckets else 0
        self.native = None
        if device.type == "cuda":
            from . import native_comm
            self.native = native_comm.SymmetricSlab(self.total, spec.n_state, device, world, rank,
                                                    buffer_floats=buffer_floats)
            self.master, self.grad, self.shadow = self.native.master, self.native.grad, self.native.shadow
            self.state = self.native.state
        else:
            self.master = torch.zeros(self.total, dtype=torch.float32, device=device)
            self.grad = torch.zeros(self.total, dtype=torch.float32, device=device)
            self.shadow = None
            self.state = [torch.zeros(self.total, dtype=torch.float32, device=device)
                          for _ in range(spec.n_state)]
        self.param_bucket: Dict[int, int] = {}
        for b in self.buckets:
            for pi in b.params:
                self.param_bucket[pi] = b.index
        self._adopt()

    @staticmethod
    def _view_like(flat: torch.Tensor, off: int, p: torch.Tensor) -> torch.Tensor:
        """A view of ``flat[off:off+numel]`` with p's logical shape AND p's physical layout
        (channels_last conv weights stay [Cout, R, S, Cin] in memory)."""
        chunk = flat[off: off + p.numel()]
        if p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last) and not p.is_contiguous():
            co, ci, r, s = p.shape
            return chunk.view(co, r, s, ci).permute(0, 3, 1, 2)
        return chunk.view(p.shape)

    def _adopt(self):
        with torch.no_grad():
            for i, p in enumerate(self.params):
                off = self.offsets[i]
                mv = self._view_like(self.master, off, p)
                mv.copy_(p.data)
                p.data = mv
                p.grad = self._view_like(self.grad, off, p)
                p._arena_index = i
                if self.shadow is not None:
                    p._shadow = self._view_like(self.shadow, off, p)
            self.refresh_shadow()

    def refresh_shadow(self):
        """Recompute the bf16 compute copy from the fp32 masters (after init / load_state_dict)."""
        if self.shadow is not None:
            self.shadow.copy_(self.master)

    def zero_grad(self):
        self.grad.zero_()


# ----------------------------------------------------------------------------- optimizer facade
class FusedOptimizer:
    """The object ``train_epoch`` calls ``zero_grad()`` / ``step()`` on (train.py:71-73).

    With the fused engine the update already happened bucket by bucket during backward, so
    ``step()`` only joins the side stream (and launches any bucket that did not fire), and
    ``zero_grad()`` is a no-op because the update kernel leaves the gradients zeroed."""

    def __init__(self, engine: "FusedDataParallel"):
        self.engine, self.spec = engine, engine.spec
        self.step_count = 0

    @property
    def param_groups(self):
        return [{"lr": self.spec.lr, "params": self.engine.arena.params}]

    def zero_grad(self, set_to_none: bool = False):
        return None

    def step(self):
        self.engine.finish_step()
        self.step_count += 1

    def state_dict(self):
        """COLLECTIVE when world > 1 on CUDA: in the two-shot / NVLS bucket kernels only the slice owner ever writes the
        momentum / Adam moments of its 1/W slice (csrc/comm/comm.cu), so the complete state is assembled by all-gathering
        the owners' slices bucket by bucket.  Every rank must call it; every rank gets the full tensors."""
        a = self.engine.arena
        return {"kind": self.spec.kind, "lr": self.spec.lr, "step": self.engine.step_index,
                "state": [s.detach().cpu() for s in self.engine.gather_sharded(a.state)]}

    def load_state_dict(self, sd):
        a = self.engine.arena
        self.spec.lr = sd["lr"]
        self.engine.step_index = sd["step"]
        for dst, src in zip(a.state, sd["state"]):
            dst.copy_(src)            # every rank keeps the full tensor; a slice owner only ever reads its own part


# ----------------------------------------------------------------------------- the engine
class FusedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, spec: Optional[OptimizerSpec] = None, process_group=None,
                 broadcast_buffers: bool = True, first_bucket_bytes=FIRST_BUCKET_BYTES,
                 bucket_cap_bytes=BUCKET_CAP_BYTES, overlap: bool = True, comm: str = "auto"):
        super().__init__()
        self.module = module
        self.spec = spec or OptimizerSpec()
        self.group = process_group
        self.world = dist.get_world_size(process_group) if is_dist() else 1
        self.rank = dist.get_rank(process_group) if is_dist() else 0
        self.overlap, self.comm = overlap, comm
        self.step_index = 0
        self._masters_stale = False     # sharded masters: a step ran since the last sync_masters()
        self.capturing = False          # True while a CUDA graph of the step is being recorded (engine/graphed.py)
        try:
            device = next(module.parameters()).device
        except StopIteration:
            device = torch.device("cpu")
        self.device = device
        self._verify_replicas()
        self._float_buffers = [b for b in module.buffers() if b.is_floating_point()]
        flat_ok = device.type == "cuda" and self._float_buffers and all(b.dtype == torch.float32 for b in self._float_buffers)
        nbuf = sum(b.numel() for b in self._float_buffers) if flat_ok else 0
        self.arena = ParamArena(module.parameters(), device, self.spec, first_bucket_bytes,
                                bucket_cap_bytes, self.world, self.rank, buffer_floats=nbuf)
        self.optimizer = FusedOptimizer(self)
        self._flat_buffers = None
        if flat_ok:
            # BN running statistics live in ONE flat region of the peer-mapped slab (the registered buffers become views of
            # it): the per-forward buffer broadcast of DDP is then a plain copy of rank 0's region into the
            # peers', done by the step's last bucket kernel between its two flag barriers -- no collective call per step
            flat = self.arena.native.bufs[:nbuf]
            off = 0
            with torch.no_grad():
                for b in self._float_buffers:
                    n = b.numel()
                    flat[off: off + n].copy_(b.reshape(-1))
                    b.data = flat[off: off + n].view(b.shape)
                    off += n
            self._flat_buffers = flat
        has_sync_bn = any(getattr(m, "sync_group", False) is not False for m in module.modules())
        # P17: with SyncBN the running stats are identical on every rank already
        self.broadcast_buffers = broadcast_buffers and self.world > 1 and not has_sync_bn
        self._hooks = []
        self._in_backward = False
        self._sync_enabled = True
        self._broadcast_initial_state()
        for i, p in enumerate(self.arena.params):
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
            p._grad_sink = self                    # native ops write wgrad straight into the arena
        self._reset_pending()
        if device.type == "cuda":
            from . import native_comm
            self._impl = native_comm.NativeReducer(self)
        else:
            self._impl = _TorchReducer(self)

    # -- replica consistency (DDP ctor: _verify_param_shape_across_processes) --
    def _verify_replicas(self):
        """Every rank must wrap the same parameter list (count, shapes, dtypes, trainability) with the same update rule:
        the arena offsets, the bucket plan and the peer-pointer arithmetic of the bucket kernels are derived from it, so
        a mismatch would corrupt peers' memory or hang the first collective.  torch DDP raises in its constructor
        (torch/nn/parallel/distributed.py:881, reached from train.py:128); so does this, before anything is allocated."""
        if self.world == 1:
            return
        import hashlib
        h = hashlib.sha1()
        n = 0
        for p in self.module.parameters():
            if p.requires_grad:
                h.update(repr((tuple(p.shape), str(p.dtype))).encode())
                n += 1
        mine = (n, h.hexdigest(), self.spec.kind)
        box = [None] * self.world
        dist.all_gather_object(box, mine, group=self.group)
        bad = [r for r, v in enumerate(box) if v != box[0]]
        if bad:
            raise RuntimeError(
                f"picd_b200: replicas disagree on the model: rank 0 wraps {box[0][0]} trainable parameters "
                f"(shape digest {box[0][1][:12]}, optimizer {box[0][2]}), rank(s) {bad} differ "
                f"(e.g. rank {bad[0]}: {box[bad[0]][0]} parameters, digest {box[bad[0]][1][:12]}, optimizer "
                f"{box[bad[0]][2]}).  All ranks must build the same network before FusedDataParallel wraps it.")

    @property
    def masters_sharded(self) -> bool:
        impl = getattr(self, "_impl", None)
        return bool(self.world > 1 and impl is not None and getattr(getattr(impl, "comm", None), "shard_master", False))

    def sync_masters(self):
        """COLLECTIVE.  With sharded masters (peer-memory engine, bf16 wire format) a rank holds the current fp32 master
        only for the bucket slices it owns -- the bf16 compute copies are what the step keeps identical everywhere.  Before
        anything reads ``parameter.data`` across the whole model (``state_dict()`` for a checkpoint, a comparison, a switch
        to the library-op path) every rank calls this: the owners' slices are all-gathered in place."""
        if not self.masters_sharded or not self._masters_stale:
            return
        impl = self._impl
        torch.cuda.current_stream(self.device).wait_stream(impl.side)
        for b in self.arena.buckets:
            if impl._mode(b) < 2:
                continue
            per = b.numel // self.world
            lo = b.start + per * self.rank
            dist.all_gather_into_tensor(self.arena.master[b.start: b.end], self.arena.master[lo: lo + per].clone(),
                                        group=self.group)
        self._masters_stale = False

    def state_dict(self, *args, **kw):
        if self.masters_sharded and self._masters_stale:
            raise RuntimeError("picd_b200: the fp32 masters are sharded by slice owner and have changed since the last "
                               "gather; call model.sync_masters() on EVERY rank before state_dict()")
        return super().state_dict(*args, **kw)

    def gather_sharded(self, flats):
        """Full copies of per-element buffers that the bucket kernels keep SHARDED by slice owner (optimizer state): for
        every bucket reduced in two-shot / NVLS mode rank r holds the valid values of slice r only.  Collective."""
        outs = [f.detach().clone() for f in flats]
        impl = getattr(self, "_impl", None)
        if self.world == 1 or self.device.type != "cuda" or impl is None or not hasattr(impl, "_mode"):
            return outs
        for b in self.arena.buckets:
            if impl._mode(b) < 2:
                continue                      # one-shot buckets: every rank updates the whole bucket
            per = b.numel // self.world
            lo = b.start + per * self.rank
            for f, o in zip(flats, outs):
                dist.all_gather_into_tensor(o[b.start: b.end], f[lo: lo + per].contiguous(), group=self.group)
        return outs

    # -- initial replica sync (DDP ctor: _sync_module_states) ------------------
    def _broadcast_initial_state(self):
        if self.world == 1:
            return
        dist.broadcast(self.arena.master, src=dist.get_global_rank(self.group, 0) if self.group else 0,
                       group=self.group)
        self.arena.refresh_shadow()
        for b in self.module.buffers():
            dist.broadcast(b, src=0, group=self.group)

    def _sync_buffers(self):
        """Per-forward coalesced broadcast of BN running stats from rank 0 -- CPU / gloo path only; on
        CUDA the step's last bucket kernel pushes rank 0's buffers over NVLink (csrc/comm/comm.cu, ``bcast_n``)."""
        if not self._float_buffers:
            return
        flat = torch.cat([b.reshape(-1).float() for b in self._float_buffers])
        dist.broadcast(flat, src=0, group=self.group)
        off = 0
        for b in self._float_buffers:
            n = b.numel()
            b.copy_(flat[off: off + n].view_as(b))
            off += n

    # -- gradient-ready tracking ---------------------------------------------------------------
    def _reset_pending(self):
        for b in self.arena.buckets:
            b.pending, b.launched, b.work = len(b.params), False, None

    def _make_hook(self, i):
        def hook(_p):
            self.mark_ready(i)
        return hook

    def mark_ready(self, param_index: int):
        """Called when parameter ``param_index``'s gradient is complete in the arena (from the
        autograd post-accumulate hook, or directly by a native wgrad op)."""
        b = self.arena.buckets[self.arena.param_bucket[param_index]]
        b.pending -= 1
        if b.pending == 0 and self._sync_enabled and self.overlap:
            self._launch(b)

    def _launch(self, b: Bucket):
        if not b.launched:
            b.launched = True
            self._impl.launch_bucket(b)

    @contextlib.contextmanager
    def no_sync(self):
        """Accumulate gradients locally (no reduction, no update) inside the context."""
        old, self._sync_enabled = self._sync_enabled, False
        try:
            yield
        finally:
            self._sync_enabled = old
            self._reset_pending()

    def finish_step(self):
        """``optimizer.step()``: launch whatever did not fire from hooks, wait, advance."""
        for b in self.arena.buckets:
            self._launch(b)
        self._impl.finish()
        self._reset_pending()
        self.step_index += 1
        self._masters_stale = True

    # -- nn.Module plumbing -------------------------------------------------------------------
    def forward(self, *args, **kw):
        if self.broadcast_buffers and self._flat_buffers is None and self.module.training and torch.is_grad_enabled() \
                and not self.capturing:
            self._sync_buffers()        # CPU / gloo (and CUDA modules with non-fp32 buffers): a collective per forward
        if self.device.type == "cuda" and self.module.training and torch.is_grad_enabled():
            from ..ops import native
            native.stat_arena(self.device).begin_step()      # one memset instead of two fill kernels per BN layer
        out = self.module(*args, **kw)
        from ..models.layers import flush_batch_counters
        flush_batch_counters()          # deferred BatchNorm num_batches_tracked updates: one multi-tensor add
        return out

    def load_state_dict(self, state_dict, strict: bool = True, **kw):
        out = super().load_state_dict(state_dict, strict=strict, **kw)
        self.arena.refresh_shadow()
        self._masters_stale = False      # every rank now holds the full, current masters
        return out


class _TorchReducer:
    """CPU / gloo execution of the bucket plan: async all-reduce per bucket, then the oracle
    update rule on the flat views.  Also the numerical reference for the CUDA kernel."""

    def __init__(self, engine: FusedDataParallel):
        self.e = engine

    def launch_bucket(self, b: Bucket):
        e = self.e
        if e.world > 1:
            b.work = dist.all_reduce(e.arena.grad[b.start: b.end], group=e.group, async_op=True)

    def finish(self):
        e, a, s = self.e, self.e.arena, self.e.spec
        step = e.step_index + 1
        with torch.no_grad():
            for b in a.buckets:
                if b.work is not None:
                    b.work.wait()
                sl = slice(b.start, b.end)
                p, g = a.master[sl], a.grad[sl]
                if s.kind == "sgd":
                    ref.sgd_momentum_step(p, g, a.state[0][sl], s.lr, s.momentum, s.weight_decay,
                                          s.nesterov, 1.0 / e.world)
                else:
                    ref.adam_step(p, g, a.state[0][sl], a.state[1][sl], step, s.lr, s.betas[0],
                                  s.betas[1], s.eps, s.weight_decay, s.kind == "adamw", 1.0 / e.world)
                g.zero_()
