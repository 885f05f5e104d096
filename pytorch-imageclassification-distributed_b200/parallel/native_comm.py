"""NVLink peer-memory backend of the data-parallel engine (CUDA only).

``SymmetricSlab``  one ``cudaMalloc`` allocation per rank holding
                   [gradient arena fp32 | master weights fp32 | bf16 shadow | exchange slots | flag words],
                   exported with CUDA IPC and mapped by every peer of the node (handles travel over the
                   existing torch.distributed group once, at construction).
                   Where torch's symmetric-memory allocator is available the slab comes from it instead, which adds an
                   NVSwitch multicast (NVLS) mapping of the same memory (``PICD_SYMM=0`` / ``PICD_NVLS=0`` opt out).
``NativeReducer``  launches ``picd_fused_allreduce_update`` per bucket on a side stream as soon as the bucket's
                   last gradient kernel has been queued (reference: DDP Reducer + NCCL + optimizer.step,
                   train.py:72-73).
``Comm``           small latency-bound collectives for SyncBatchNorm over the same slab
                   (reference: torch/nn/modules/_functions.py:74-76,158-164).
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..ops._lib import check, require_lib

SLOT_FLOATS = 2 * 2048 + 8          # largest payload: (sum, sumsq) for up to 2048 channels + count
N_SLOTS = 1024                      # small-collective slots (flag words); eager and graph-captured launches use disjoint
                                    # slots, so a ResNet-101 needs 4 x 104 of them
XCHG_FLOATS = 16 << 20              # exchange region (64 MiB): a slot takes 2 halves x world sub-slots x its payload, handed
                                    # out by a bump allocator (ResNet-101 at 8 ranks: 4 x 2 x 8 x 105k floats = 27 MiB)
MAX_PEERS = 8
TWO_SHOT_MIN_BYTES = 256 << 10      # below this a bucket is latency-bound: one-shot
NVLS_MIN_BYTES = 4 << 20            # from here on the in-switch reduction (multimem.ld_reduce) replaces the W peer loads, if mapped
NVLS_ST_MAX_BYTES = 512 << 20       # above this the multimem.st broadcast is slower than unicast pushes (measured at W=8): mode 4
NVLS_MIN_WORLD = 4                  # with 2 ranks a multicast access moves as many bytes as a peer access and is slower per SM

_current: Optional["Comm"] = None


def current_comm() -> Optional["Comm"]:
    return _current


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


class _DevMem:
    """Expose a raw device range through __cuda_array_interface__ so torch can view it without copying."""

    def __init__(self, ptr: int, shape, typestr: str, owner):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}
        self._owner = owner


def _view(ptr: int, n: int, typestr: str, device, owner) -> torch.Tensor:
    return torch.as_tensor(_DevMem(ptr, (n,), typestr, owner), device=device)


def pass_fds(rank: int, world: int, my_fd: int, owners, device=None) -> dict:
    """Move POSIX file descriptors between the ranks of this node: for every `owner` rank in turn, the owner serves its
    descriptor over a unix socket (SCM_RIGHTS) and every other rank connects and receives a descriptor of ITS OWN
    process that refers to the same kernel object.  Returns {owner rank: local fd}.  Collective."""
    import socket
    import tempfile
    import time
    token = [None]
    if rank == 0:
        token[0] = f"{os.getpid()}_{int(time.time() * 1e6) & 0xFFFFFFFF:x}"
    dist.broadcast_object_list(token, src=0)
    got = {}
    failed = False

    def barrier():
        dist.barrier(device_ids=[device.index]) if (device is not None and dist.get_backend() == "nccl") else dist.barrier()

    for owner in owners:
        path = os.path.join(tempfile.gettempdir(), f"picd_fd_{token[0]}_{owner}.sock")
        srv = None
        if rank == owner:
            try:
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                srv.settimeout(30)
                srv.bind(path)
                srv.listen(world)
            except OSError:
                failed, srv = True, None
        barrier()                                        # the socket exists (or its owner has given up): peers may connect
        try:                                             # both barriers run on every rank whatever happens in between
            if rank == owner:
                if srv is not None:
                    for _ in range(world - 1):
                        conn, _ = srv.accept()
                        socket.send_fds(conn, [b"f"], [my_fd])
                        conn.close()
            else:
                c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                c.settimeout(30)
                c.connect(path)
                _, fds, _, _ = socket.recv_fds(c, 1, 1)
                c.close()
                got[owner] = fds[0]
        except OSError:
            failed = True
        barrier()
        if srv is not None:
            srv.close()
            try:
                os.unlink(path)
            except OSError:
                pass
    if failed:
        for f in got.values():
            os.close(f)
        raise OSError("picd_b200: file-descriptor exchange over unix sockets failed")
    return got


class SymmetricSlab:
    def __init__(self, total: int, n_state: int, device: torch.device, world: int, rank: int,
                 num_buckets_hint: int = 256, buffer_floats: int = 0):
        self.lib = require_lib()
        self.total, self.device, self.world, self.rank = total, device, world, rank
        if world > MAX_PEERS:
            raise RuntimeError(f"picd_b200: the peer-memory engine supports up to {MAX_PEERS} ranks of one node "
                               f"(got world_size={world})")
        self.off_grad = 0
        self.off_master = _align(self.off_grad + total * 4)
        self.off_shadow = _align(self.off_master + total * 4)
        self.off_xchg = _align(self.off_shadow + total * 2)
        self.off_flags = _align(self.off_xchg + XCHG_FLOATS * 4)
        self.flag_words = (2 * num_buckets_hint + N_SLOTS) * MAX_PEERS
        # module buffers (BatchNorm running statistics): rank 0's copy is pushed into the peers' by the step's last bucket kernel
        self.buffer_floats = (int(buffer_floats) + 3) // 4 * 4
        self.off_bufs = _align(self.off_flags + self.flag_words * 4)
        # bf16 staging copy of the gradient arena: the wire format of the reduction (peers read 2 B per element, not 4)
        self.off_gbf = _align(self.off_bufs + self.buffer_floats * 4)
        self.nbytes = _align(self.off_gbf + (total * 2 if world > 1 else 0))
        self.max_buckets = num_buckets_hint
        self.mc_base = 0          # NVSwitch multicast (NVLS) view of the slab, when the host can map one
        self._symm = None         # torch symmetric-memory allocation, if that is where the slab came from
        self._vmm = None          # handle of the library's own VMM slab (csrc/comm/vmm_slab.cu), if that is
        self.backing = "cudaMalloc+IPC"
        mode = os.environ.get("PICD_SYMM", "own")
        if world > 1 and mode not in ("0", "torch"):
            self._try_vmm_alloc()             # own path: cuMemCreate + fd exchange + cuMulticast*
        if world > 1 and self._vmm is None and mode != "0":
            self._try_symmetric_alloc()       # torch's symmetric-memory allocator
        if self._symm is None and self._vmm is None:
            base = ctypes.c_void_p()
            with torch.cuda.device(device):
                self.lib.picd_slab_alloc.argtypes = [ctypes.c_size_t, ctypes.POINTER(ctypes.c_void_p)]
                check(self.lib.picd_slab_alloc(self.nbytes, ctypes.byref(base)), "picd_slab_alloc")
            self.base = base.value
        self.grad = _view(self.base + self.off_grad, total, "<f4", device, self)
        self.master = _view(self.base + self.off_master, total, "<f4", device, self)
        self.shadow = _view(self.base + self.off_shadow, total, "<i2", device, self).view(torch.bfloat16)
        self.state = [torch.zeros(total, dtype=torch.float32, device=device) for _ in range(n_state)]
        self.bufs = _view(self.base + self.off_bufs, self.buffer_floats, "<f4", device, self) if self.buffer_floats else None
        if self._symm is None and self._vmm is None:
            self.peer_bases: List[int] = [self.base] * max(world, 1)
            if world > 1:
                self._exchange_handles()

    # -- own VMM slab -------------------------------------------------------------------------------------------
    def _agree(self, ok: bool) -> bool:
        """True only if EVERY rank reports success (collective), so that all ranks take the same path."""
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        return int(flag.item()) == 1

    def _pass_fds(self, my_fd: int, owners) -> dict:
        return pass_fds(self.rank, self.world, my_fd, owners, self.device)

    def _try_vmm_alloc(self):
        """The library's own symmetric slab: a VMM allocation per rank exported as a POSIX fd, every peer's allocation mapped
        into every process, and -- where the GPUs support it -- an NVSwitch multicast object over all of them whose mapping
        is what the NVLS variant of the bucket kernel uses (multimem.ld_reduce / multimem.st).  Every step is followed by
        an agreement round, so a failure on any rank sends ALL ranks on to the next allocator."""
        lib = self.lib
        if not hasattr(lib, "picd_vmm_create"):
            return
        lib.picd_vmm_create.argtypes = [ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p),
                                        ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int),
                                        ctypes.POINTER(ctypes.c_size_t)]
        lib.picd_vmm_import_peer.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
        lib.picd_vmm_mc_create.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        lib.picd_vmm_mc_import.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.picd_vmm_mc_add_device.argtypes = [ctypes.c_void_p]
        lib.picd_vmm_mc_bind_map.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
        lib.picd_vmm_destroy.argtypes = [ctypes.c_void_p]
        slab, base, fd, padded = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int(-1), ctypes.c_size_t(0)
        with torch.cuda.device(self.device):
            ok = lib.picd_vmm_create(self.nbytes, self.world, ctypes.byref(slab), ctypes.byref(base), ctypes.byref(fd),
                                     ctypes.byref(padded)) == 0
            if not self._agree(ok):
                if ok:
                    lib.picd_vmm_destroy(slab)
                return
            bases = [0] * self.world
            bases[self.rank] = base.value
            try:
                fds = self._pass_fds(fd.value, range(self.world))
                ok = True
                for r, pfd in fds.items():
                    p = ctypes.c_void_p()
                    ok = ok and lib.picd_vmm_import_peer(slab, r, pfd, ctypes.byref(p)) == 0
                    os.close(pfd)
                    bases[r] = p.value or 0
            except Exception:      # noqa: BLE001 -- no unix sockets / fd passing in this container
                ok = False
            os.close(fd.value)
            if not self._agree(ok):
                lib.picd_vmm_destroy(slab)
                return
            # multicast view (NVLS); optional: without it the slab still serves the P2P variants
            mc_base = 0
            want_mc = os.environ.get("PICD_NVLS", "1") != "0" and lib.picd_vmm_multicast_supported() == 1
            if self._agree(want_mc):
                mfd = ctypes.c_int(-1)
                ok = True
                if self.rank == 0:
                    ok = lib.picd_vmm_mc_create(slab, self.world, ctypes.byref(mfd)) == 0
                if self._agree(ok):
                    try:
                        got = self._pass_fds(mfd.value, [0])
                        if self.rank != 0:
                            ok = lib.picd_vmm_mc_import(slab, got[0]) == 0
                            os.close(got[0])
                        else:
                            os.close(mfd.value)
                    except Exception:      # noqa: BLE001
                        ok = False
                    ok = ok and lib.picd_vmm_mc_add_device(slab) == 0
                    if self._agree(ok):            # every device is a member before anybody binds
                        p = ctypes.c_void_p()
                        ok = lib.picd_vmm_mc_bind_map(slab, ctypes.byref(p)) == 0
                        if self._agree(ok):
                            mc_base = p.value
            self._vmm = slab
            self.base = base.value
            self.peer_bases = bases
            self.mc_base = mc_base or 0
            self.backing = "own VMM slab" + (" + multicast" if mc_base else "")

    def _try_symmetric_alloc(self):
        """Allocate the slab from torch's symmetric-memory allocator: besides the peer mappings (what the CUDA-IPC path
        below provides) it returns an NVSwitch MULTICAST mapping of the same memory where the host supports NVLS, which
        mode 3 of the bucket kernel uses (multimem.ld_reduce / multimem.st).  Collective; every rank must agree on the
        outcome, so a failure anywhere sends all ranks to the cudaMalloc + IPC path."""
        ok, keep = 1, None
        try:
            import torch.distributed._symmetric_memory as sm
            t = sm.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            t.zero_()
            torch.cuda.synchronize(self.device)
            h = sm.rendezvous(t, dist.group.WORLD)
            bases = [int(p) for p in h.buffer_ptrs]
            if len(bases) != self.world or bases[self.rank] != t.data_ptr():
                ok = 0
            keep = (t, h, bases)
        except Exception:      # noqa: BLE001 -- older torch, no fabric support, container without fd passing, ...
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) != 1:
            return
        t, h, bases = keep
        self._symm = (t, h)
        self.base = t.data_ptr()
        self.peer_bases = bases
        self.mc_base = int(h.multicast_ptr) if os.environ.get("PICD_NVLS", "1") != "0" else 0
        self.backing = "torch symmetric memory" + (" + multicast" if self.mc_base else "")

    def _exchange_handles(self):
        nb = self.lib.picd_ipc_handle_bytes()
        buf = ctypes.create_string_buffer(nb)
        self.lib.picd_ipc_get_handle.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        check(self.lib.picd_ipc_get_handle(self.base, buf), "picd_ipc_get_handle")
        handles = [None] * self.world
        dist.all_gather_object(handles, (self.rank, bytes(buf.raw)))
        self.lib.picd_ipc_open.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
        bases = [0] * self.world
        with torch.cuda.device(self.device):
            for r, raw in handles:
                if r == self.rank:
                    bases[r] = self.base
                    continue
                p = ctypes.c_void_p()
                check(self.lib.picd_ipc_open(ctypes.create_string_buffer(raw, nb), ctypes.byref(p)),
                      f"picd_ipc_open(rank {r})")
                bases[r] = p.value
        self.peer_bases = bases
        dist.barrier(device_ids=[self.device.index]) if dist.get_backend() == "nccl" else dist.barrier()

    def region_ptrs(self, off: int):
        arr = (ctypes.c_void_p * MAX_PEERS)()
        for i, b in enumerate(self.peer_bases):
            arr[i] = b + off
        return arr


class Comm:
    """Owns the native communicator context for one slab."""

    def __init__(self, slab: SymmetricSlab, num_buckets: int):
        global _current
        self.slab, self.lib = slab, slab.lib
        if num_buckets > slab.max_buckets:
            raise RuntimeError(f"picd_b200: {num_buckets} buckets exceed the slab's flag space")
        self.num_buckets = num_buckets
        ctx = ctypes.c_void_p()
        f = self.lib.picd_comm_create
        f.argtypes = [ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 5 + [ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
        with torch.cuda.device(slab.device):
            check(f(slab.rank, max(slab.world, 1), slab.region_ptrs(slab.off_grad), slab.region_ptrs(slab.off_master),
                    slab.region_ptrs(slab.off_shadow), slab.region_ptrs(slab.off_xchg),
                    slab.region_ptrs(slab.off_flags), slab.max_buckets, ctypes.byref(ctx)), "picd_comm_create")
        self.ctx = ctx
        self.has_nvls = bool(slab.mc_base)
        if self.has_nvls:
            f2 = self.lib.picd_comm_set_multicast
            f2.argtypes = [ctypes.c_void_p] * 4
            check(f2(ctx, slab.mc_base + slab.off_grad, slab.mc_base + slab.off_master, slab.mc_base + slab.off_shadow),
                  "picd_comm_set_multicast")
        if slab.buffer_floats:
            f3 = self.lib.picd_comm_set_buffers
            f3.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong]
            check(f3(ctx, slab.region_ptrs(slab.off_bufs), slab.buffer_floats), "picd_comm_set_buffers")
        # wire format of the gradient exchange (PICD_GRAD_WIRE=fp32 restores full-precision exchange + replicated masters)
        self.wire_bf16 = slab.world > 1 and os.environ.get("PICD_GRAD_WIRE", "bf16") != "fp32"
        # Sharded masters (opt-in, PICD_SHARD_MASTERS=1): the slice owner keeps its fp32 master slice to itself and pushes only
        # the bf16 compute copy -- 2 B per element on the wire instead of 6.  Safe only when EVERY forward read of a parameter
        # goes through the native ops (which read the bf16 copies): a torch op applied to a parameter directly (nn.Linear,
        # a position embedding added with `+`) reads the fp32 master, which is stale on the ranks that do not own its slice.
        # Hence off by default, and never with the library-op A/B switch.
        self.shard_master = self.wire_bf16 and os.environ.get("PICD_SHARD_MASTERS", "0") == "1" and \
            os.environ.get("PICD_FORCE_TORCH", "0") != "1"
        if slab.world > 1:
            f5 = self.lib.picd_comm_set_wire
            f5.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
            check(f5(ctx, slab.region_ptrs(slab.off_gbf), (slab.mc_base + slab.off_gbf) if slab.mc_base else None,
                     int(self.wire_bf16), int(self.shard_master)), "picd_comm_set_wire")
        timeout_s = float(os.environ.get("PICD_FLAG_TIMEOUT_S", "600"))
        f4 = self.lib.picd_comm_set_flag_timeout
        f4.argtypes = [ctypes.c_double]
        with torch.cuda.device(slab.device):
            check(f4(timeout_s), "picd_comm_set_flag_timeout")
        self._slot_epoch = {}
        self._slots = {}
        self._slot_keys = []
        self._next_slot = 0
        self._bump = 0                 # floats of the exchange region handed out so far
        self.graph_dyn = None          # device float[8] while a CUDA graph of the step is recorded (word 4 = replay delta)
        self._graph_slots = []         # slots owned by captured launches: their epochs advance once per replay
        self._f_upd = self.lib.picd_fused_allreduce_update
        self._f_upd.argtypes = ([ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, ctypes.c_longlong,
                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_float] * 8 +
                                [ctypes.c_int, ctypes.c_uint, ctypes.c_int, ctypes.c_int, ctypes.c_void_p])
        self._f_upd_dyn = self.lib.picd_fused_allreduce_update_dyn
        self._f_upd_dyn.argtypes = ([ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, ctypes.c_longlong,
                                     ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_float] * 5 +
                                    [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p])
        self._f_adv = self.lib.picd_dyn_advance
        self._f_adv.argtypes = [ctypes.c_void_p, ctypes.c_float, ctypes.c_float, ctypes.c_void_p]
        self._f_small_dyn = self.lib.picd_p2p_allreduce_small_dyn
        self._f_small_dyn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                      ctypes.c_longlong, ctypes.c_int, ctypes.c_uint, ctypes.c_void_p, ctypes.c_int,
                                      ctypes.c_int, ctypes.c_void_p]
        self._f_bn_dyn = self.lib.picd_bn_finalize_sync_dyn
        self._f_bn_dyn.argtypes = ([ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_float] + [ctypes.c_void_p] * 4 +
                                   [ctypes.c_float, ctypes.c_float] + [ctypes.c_void_p] * 5 +
                                   [ctypes.c_int, ctypes.c_longlong, ctypes.c_int, ctypes.c_uint, ctypes.c_void_p, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_void_p])
        self._f_small = self.lib.picd_p2p_allreduce_small
        self._f_small.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                  ctypes.c_longlong, ctypes.c_int, ctypes.c_uint, ctypes.c_int, ctypes.c_void_p]
        self._f_bn = self.lib.picd_bn_finalize_sync
        self._f_bn.argtypes = ([ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_float] + [ctypes.c_void_p] * 4 +
                               [ctypes.c_float, ctypes.c_float] + [ctypes.c_void_p] * 5 +
                               [ctypes.c_int, ctypes.c_longlong, ctypes.c_int, ctypes.c_uint, ctypes.c_int, ctypes.c_void_p])
        _current = self

    # -- gradient buckets ---------------------------------------------------------------------------
    def fused_update(self, bucket_index: int, mode: int, start: int, count: int, state, spec, step: int,
                     epoch: int, blocks: int = 0, dyn: Optional[torch.Tensor] = None, bcast_buffers: bool = False):
        s0 = state[0].data_ptr()
        s1 = state[1].data_ptr() if len(state) > 1 else None
        b1, b2 = spec.betas
        if dyn is not None:      # CUDA-graph capture: lr / bias corrections / epoch are read from `dyn` at run time
            check(self._f_upd_dyn(self.ctx, bucket_index, mode, start, count, s0, s1, spec.KINDS[spec.kind],
                                  spec.momentum, b1, b2, spec.eps, spec.weight_decay, int(spec.nesterov), blocks,
                                  dyn.data_ptr(), int(bcast_buffers), torch.cuda.current_stream().cuda_stream),
                  "picd_fused_allreduce_update_dyn")
            return
        bc1, bc2 = 1.0 - b1 ** step, 1.0 - b2 ** step
        check(self._f_upd(self.ctx, bucket_index, mode, start, count, s0, s1, spec.KINDS[spec.kind], spec.lr,
                          spec.momentum, b1, b2, spec.eps, spec.weight_decay, bc1, bc2, int(spec.nesterov),
                          epoch & 0xFFFFFFFF, blocks, int(bcast_buffers), torch.cuda.current_stream().cuda_stream),
              "picd_fused_allreduce_update")

    # -- small collectives (SyncBN) ------------------------------------------------------------------
    def _slot(self, key, n: int):
        """(slot index, float offset of its even half, sub-slot floats) of the exchange slot of `key`; sized at first use for a
        payload of n floats: 2 halves (epoch parity) x world sub-slots (one per pushing rank).  Every rank calls this in the
        same order, so the layout agrees across the node."""
        k = (id(key), self.graph_dyn is not None)        # captured launches get slots of their own
        s = self._slots.get(k)
        if s is None:
            sub = (n + 3) // 4 * 4
            need = 2 * max(self.slab.world, 1) * sub
            if self._next_slot >= N_SLOTS or self._bump + need > XCHG_FLOATS:
                raise RuntimeError("picd_b200: out of small-collective slots")
            s = (self._next_slot, self._bump, sub)
            self._next_slot += 1
            self._bump += need
            self._slots[k] = s
            self._slot_keys.append(key)          # keep the key alive so its id() stays unique
        if s[2] < n:
            raise RuntimeError("picd_b200: small-collective slot reused with a larger payload")
        return s

    def _next(self, slot):
        index, base, sub = slot
        half = max(self.slab.world, 1) * sub
        e = self._slot_epoch.get(index, 0) + 1
        self._slot_epoch[index] = e
        fslot = 2 * self.slab.max_buckets + index
        if self.graph_dyn is not None:
            # recorded launch: pass the EVEN half and the capture-time epoch; the kernel adds the replay delta itself
            self._graph_slots.append(index)
            return e, base, fslot, half, sub
        return e, base + (e & 1) * half, fslot, half, sub

    def begin_capture(self, dyn: torch.Tensor):
        self.graph_dyn, self._graph_slots = dyn, []

    def end_capture(self):
        self.graph_dyn = None

    def note_replay(self, index: int):
        """Host bookkeeping for replay number `index` (0 = first): the graph's slots move one epoch per replay."""
        if index > 0:
            for s in self._graph_slots:
                self._slot_epoch[s] += 1

    def _delta_ptr(self):
        return self.graph_dyn.data_ptr() + 16          # word 4 of the dyn array

    def bn_finalize_sync(self, sums, count, gamma, beta, rm, rv, momentum, eps, coef) -> float:
        C = gamma.numel()
        if 2 * C + 1 > SLOT_FLOATS:
            raise RuntimeError(f"picd_b200: SyncBN with {C} channels exceeds the exchange slot")
        e, off, fslot, half, sub = self._next(self._slot(gamma, 2 * C + 1))
        total = torch.empty(1, dtype=torch.float32, device=sums.device)
        args = (self.ctx, sums.data_ptr(), sums.numel() // (2 * C), float(count), gamma.data_ptr(), beta.data_ptr(),
                None if rm is None else rm.data_ptr(), None if rv is None else rv.data_ptr(),
                float(momentum), float(eps), coef[0].data_ptr(), coef[1].data_ptr(), coef[2].data_ptr(),
                coef[3].data_ptr(), total.data_ptr(), C, off, fslot, e & 0xFFFFFFFF)
        st = torch.cuda.current_stream().cuda_stream
        if self.graph_dyn is not None:
            check(self._f_bn_dyn(*args, self._delta_ptr(), half, sub, st), "picd_bn_finalize_sync_dyn")
        else:
            check(self._f_bn(*args, sub, st), "picd_bn_finalize_sync")
        # `total` (device) is the exchanged GLOBAL count: the backward scales with it, so ranks may hold different batch
        # sizes (a ragged last batch through a non-padding loader); the host-side float is only an estimate for callers
        # that want a number without a sync
        return float(count) * self.slab.world, total

    def allreduce_small(self, vec: torch.Tensor, n: Optional[int] = None, key=None) -> torch.Tensor:
        """Sum ``n`` floats over all ranks.  ``vec`` may hold several replicas of the payload ([R, n]); they are
        folded while publishing.  Returns a single [n] tensor."""
        n = vec.numel() if n is None else n
        replicas = vec.numel() // n
        if n > SLOT_FLOATS:
            raise RuntimeError("picd_b200: small all-reduce payload exceeds the exchange slot")
        holder = _BwdKey.of(key) if key is not None else _BwdKey.ring(n)
        e, off, fslot, half, sub = self._next(self._slot(holder, n if key is not None else SLOT_FLOATS))
        out = torch.empty(n, dtype=vec.dtype, device=vec.device)
        st = torch.cuda.current_stream().cuda_stream
        if self.graph_dyn is not None:
            check(self._f_small_dyn(self.ctx, vec.data_ptr(), replicas, out.data_ptr(), n, off, fslot, e & 0xFFFFFFFF,
                                    self._delta_ptr(), half, sub, st), "picd_p2p_allreduce_small_dyn")
        else:
            check(self._f_small(self.ctx, vec.data_ptr(), replicas, out.data_ptr(), n, off, fslot, e & 0xFFFFFFFF, sub, st),
                  "picd_p2p_allreduce_small")
        return out


class _BwdKey:
    """Slot identity of a backward reduction.  With a `key` (the layer's gamma parameter) every layer owns a slot sized for
    its payload; without one the caller gets a small ring of full-size slots indexed by call order (every rank runs the
    same sequence of calls, so positions agree across ranks)."""
    _by_key = {}
    _ring: List["_BwdKey"] = []
    _cursor = 0
    RING = 16

    @classmethod
    def of(cls, key):
        k = cls._by_key.get(id(key))
        if k is None:
            k = cls._by_key[id(key)] = cls()
            k.anchor = key                # keeps id(key) unique for the life of the process
        return k

    @classmethod
    def ring(cls, n):
        if not cls._ring:
            cls._ring = [cls() for _ in range(cls.RING)]
        k = cls._ring[cls._cursor % cls.RING]
        cls._cursor += 1
        return k


class NativeReducer:
    def __init__(self, engine):
        self.e = engine
        a = engine.arena
        self.comm = Comm(a.native, len(a.buckets))
        self.side = torch.cuda.Stream(device=engine.device)
        self._epoch = 0
        self._dirty = False
        self._launched = 0         # buckets launched so far in the current step
        self.dyn = None            # device float[8] (lr, bc1, bc2, epoch bits, replay delta bits) while a step is captured / replayed

    def enable_dyn(self):
        if self.dyn is None:
            self.dyn = torch.zeros(8, dtype=torch.float32, device=self.e.device)
        return self.dyn

    def advance_dyn(self):
        """Recorded as the FIRST kernel of a captured step: the device itself moves the step counter / flag epoch, the
        replay index and the Adam bias corrections forward (``picd_dyn_advance``), so a replay needs no H2D copy."""
        b1, b2 = self.e.spec.betas
        check(self.comm._f_adv(self.dyn.data_ptr(), b1, b2, torch.cuda.current_stream().cuda_stream), "picd_dyn_advance")

    def sync_dyn(self, step: int, replay_index: int = 0):
        """Make sure the captured kernels of the coming replay see (lr, step, replay index): the device advances step and
        replay index by itself, so the host only writes when its own bookkeeping says the device copy is off -- the first
        replay, after an eagerly executed step in between (ragged batch), or when the scheduler changed the learning rate.
        The write is a small synchronous copy of the values ONE BEFORE the wanted ones (the graph's first kernel adds one)."""
        s = self.e.spec
        want = (step & 0xFFFFFFFF, replay_index & 0xFFFFFFFF, float(s.lr))
        if getattr(self, "_dyn_next", None) != want:
            b1, b2 = s.betas
            host = np.zeros(8, dtype=np.float32)
            host[0] = s.lr
            prev = max(step - 1, 0)
            host[1], host[2] = 1.0 - b1 ** prev, 1.0 - b2 ** prev
            host.view(np.uint32)[3] = (step - 1) & 0xFFFFFFFF       # raw bits (never touched as floats)
            host.view(np.uint32)[4] = (replay_index - 1) & 0xFFFFFFFF
            self.dyn.copy_(torch.from_numpy(host))
        self._dyn_next = ((step + 1) & 0xFFFFFFFF, (replay_index + 1) & 0xFFFFFFFF, float(s.lr))
        self.comm.note_replay(replay_index)

    def _mode(self, b) -> int:
        if self.e.world == 1:
            return 0
        forced = self.e.comm
        if forced == "one_shot":
            return 1
        if forced == "two_shot":
            return 2
        if forced in ("nvls", "nvls_ld"):
            if not self.comm.has_nvls:
                raise RuntimeError("picd_b200: comm='nvls' but no NVSwitch multicast mapping is available")
            return 3 if forced == "nvls" else 4
        if b.numel * 4 < TWO_SHOT_MIN_BYTES:
            return 1
        if self.comm.has_nvls and self.e.world >= NVLS_MIN_WORLD and b.numel * 4 >= NVLS_MIN_BYTES:
            return 3 if b.numel * 4 < NVLS_ST_MAX_BYTES else 4
        return 2

    def launch_bucket(self, b):
        e = self.e
        cur = torch.cuda.current_stream(e.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.side.wait_event(ev)
        self._launched += 1
        # the step's last bucket also carries rank 0's module buffers to the peers (DDP's per-forward buffer broadcast)
        bcast = e.broadcast_buffers and e._flat_buffers is not None and self._launched == len(e.arena.buckets)
        with torch.cuda.stream(self.side):
            self.comm.fused_update(b.index, self._mode(b), b.start, b.numel, e.arena.state, e.spec,
                                   e.step_index + 1, e.step_index + 1, dyn=self.dyn if e.capturing else None,
                                   bcast_buffers=bcast)
        self._dirty = True

    def finish(self):
        self._launched = 0
        if self._dirty:
            torch.cuda.current_stream(self.e.device).wait_stream(self.side)
            self._dirty = False


def nvls_supported() -> int:
    """1 / 0 / -1: whether the device exposes NVSwitch multicast (NVLS).  Informational: whether the engine actually
    uses its NVLS kernel variant depends on a multicast MAPPING of the slab being available (``Comm.has_nvls``, obtained
    through torch's symmetric-memory allocator in ``SymmetricSlab._try_symmetric_alloc``)."""
    lib = require_lib()
    lib.picd_nvls_supported.restype = ctypes.c_int
    return int(lib.picd_nvls_supported())


def calibrate_two_shot_threshold(device: torch.device, sizes_bytes=(64 << 10, 256 << 10, 1 << 20, 4 << 20),
                                 iters: int = 10) -> int:
    """Measure one-shot vs two-shot on this machine and set ``TWO_SHOT_MIN_BYTES`` to the smallest bucket size at which
    two-shot wins (BASELINE.json: "the variant picked per bucket size from measured bus bandwidth").  Collective: every
    rank must call it; timings are max-reduced so all ranks agree.  Returns the chosen threshold in bytes."""
    global TWO_SHOT_MIN_BYTES
    from .engine import FusedDataParallel, OptimizerSpec
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        return TWO_SHOT_MIN_BYTES
    n = max(sizes_bytes) // 4
    holder = torch.nn.Module()
    holder.w = torch.nn.Parameter(torch.zeros(n, device=device))
    eng = FusedDataParallel(holder, OptimizerSpec("sgd", lr=0.0, momentum=0.0), overlap=False)
    comm, state, spec = eng._impl.comm, eng.arena.state, eng.spec
    epoch = 0
    chosen = None
    for size in sorted(sizes_bytes):
        count = size // 4 // (64 * world) * (64 * world)
        t = {}
        for mode in (1, 2):
            for it in range(3 + iters):
                if it == 3:
                    torch.cuda.synchronize(device)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                epoch += 1
                comm.fused_update(0, mode, 0, count, state, spec, 1, epoch, 0)
            b.record()
            torch.cuda.synchronize(device)
            ms = torch.tensor([a.elapsed_time(b) / iters], device=device)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            t[mode] = float(ms)
        if chosen is None and t[2] < t[1] * 0.98:
            chosen = size
    TWO_SHOT_MIN_BYTES = chosen if chosen is not None else max(sizes_bytes) * 4
    return TWO_SHOT_MIN_BYTES
