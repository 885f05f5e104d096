"""CUDA execution path: autograd Functions over ``libpicd_b200.so`` (C ABI in include/picd_b200.h).

Layout contract: 4-D activations are logical NCHW tensors in ``torch.channels_last`` memory
format (physically NHWC) in bf16; conv weights are channels_last too (physically
[Cout, R, S, Cin]).  Parameters are fp32 masters; the kernels read a bf16 *shadow* copy
(``param._shadow``, maintained by the fused optimizer, or cast on demand here).

Gradient routing: when a parameter belongs to a :class:`FusedDataParallel` arena
(``param._grad_sink``), wgrad / dgamma / dbeta kernels accumulate straight into the arena
view ``param.grad`` and the Function reports readiness to the engine itself
(``mark_ready``) instead of returning a gradient tensor -- no extra accumulate kernel, and
the bucket's fused all-reduce + update can start right behind the wgrad launch.
"""
from __future__ import annotations

import ctypes
import os
import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import reference as ref
from ._lib import check, require_lib

STAT_REPLICAS = 16        # must equal picd_stat_replicas(); checked at first use
ACT = {None: 0, "none": 0, "relu": 1, "gelu": 2, "swish": 3, "silu": 3, "sigmoid": 4}

_SIG = {
    "picd_conv_fprop": "pppppi" + "i" * 12 + "p",
    "picd_conv_dgrad": "ppp" + "i" * 12 + "p",
    "picd_conv_dgrad_accumulate": "ppp" + "i" * 12 + "p",
    "picd_conv_dgrad_bnreduce": "ppp" + "i" * 12 + "ppppp" + "i" + "p" + "p",
    "picd_conv_wgrad": "ppp" + "i" * 12 + "p",
    "picd_attention_fwd": "pppiiip",
    "picd_attention_bwd": "ppppppiiip",
    "picd_depthwise_fprop": "ppp" + "i" * 10 + "p",
    "picd_depthwise_dgrad": "ppp" + "i" * 10 + "p",
    "picd_depthwise_wgrad": "ppp" + "i" * 10 + "p",
    "picd_gemm_nt": "ppppiiiip",
    "picd_gemm_nn": "pppiiip",
    "picd_gemm_tn_f32": "pppiiip",
    "picd_gemm_tn_f32_bias": "ppppiiip",
    "picd_conv_wgrad_bias": "pppp" + "i" * 12 + "p",
    "picd_bn_stats": "pplip",
    "picd_bn_finalize": "pifppppffppppip",
    "picd_bn_eval_coeffs": "ppppfppip",
    "picd_bn_apply": "ppppiplip",
    "picd_bn_apply_mask": "ppppipplip",
    "picd_bn_finalize_apply": "ppifppppffpppppipplip",
    "picd_bn_bwd_reduce_mask": "ppppppppiplip",
    "picd_bn_bwd_apply_mask": "ppppppppi" + "pif" + "pi" + "pp" + "pp" + "lip",
    "picd_bn_bwd_apply_mask_cnt": "ppppppppi" + "pif" + "p" + "pi" + "pp" + "pp" + "lip",
    "picd_bn_bwd_reduce": "pppppppiplip",
    "picd_bn_bwd_apply": "pppppppi" + "pif" + "pi" + "pp" + "pp" + "lip",
    "picd_bn_bwd_eval": "pppppipplip",
    "picd_bn_relu_maxpool_fwd": "ppppp" + "iiiiii" + "p",
    "picd_bn_relu_maxpool_bwd_reduce": "pppppppp" + "iiiiii" + "p",
    "picd_bn_relu_maxpool_bwd_apply": "ppppppp" + "pifp" + "pi" + "pp" + "p" + "iiiiii" + "p",
    "picd_pool_fwd": "ppp" + "i" * 11 + "p",
    "picd_pool_bwd": "ppp" + "i" * 11 + "p",
    "picd_channel_scale_fwd": "ppppiiip",
    "picd_channel_scale_bwd": "ppppppiiip",
    "picd_gap_fwd": "ppiiip",
    "picd_gap_fwd_split": "pppiiip",
    "picd_gap_bwd": "ppiiip",
    "picd_cross_entropy": "pppppiip",
    "picd_act_fwd": "ppplii" + "p",
    "picd_act_bwd": "ppppliip",
    "picd_act_bwd_bias": "pppppliip",
    "picd_colsum": "pplip",
    "picd_dropout": "pplfup",
    "picd_dropout_dyn": "pplfupp",
    "picd_layernorm_fwd": "pppppplifp",
    "picd_layernorm_bwd": "pppppppplip",
    "picd_add_layernorm_fwd": "pppppppplifp",
    "picd_layernorm_bwd_add": "ppppppppplip",
    "picd_small_linear_fwd": "ppppiiiiip",
    "picd_small_linear_bwd": "pppppppp" + "iiiiii" + "p",
    "picd_pack_s2d": "pp" + "i" * 11 + "p",
    "picd_synthetic_batch": "ppiiiiiuip",
    "picd_normalize_u8": "pplii" + "p",
    "picd_cast_f32_bf16": "pplp",
    "picd_concat_nhwc": "ppiplip",
    "picd_top1_hits": "pippiip",
}
_CT = {"p": ctypes.c_void_p, "i": ctypes.c_int, "f": ctypes.c_float, "l": ctypes.c_longlong,
       "u": ctypes.c_uint, "z": ctypes.c_size_t}
_bound = {}


def _fn(name: str):
    f = _bound.get(name)
    if f is None:
        lib = require_lib()
        f = getattr(lib, name)
        f.argtypes = [_CT[c] for c in _SIG[name]]
        f.restype = ctypes.c_int
        _bound[name] = f
    return f


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


_sm_counts = {}


def _sm_count(device) -> int:
    d = torch.device(device)
    idx = d.index if d.index is not None else torch.cuda.current_device()
    n = _sm_counts.get(idx)
    if n is None:
        n = _sm_counts[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return n


def _stream():
    return torch.cuda.current_stream().cuda_stream


def call(name: str, *args):
    """Invoke a library entry point on the current CUDA stream (appended automatically)."""
    conv = [(_p(a) if torch.is_tensor(a) else a) for a in args]
    check(_fn(name)(*conv, _stream()), name)


# --------------------------------------------------------------------------- helpers
def nhwc(x: torch.Tensor) -> torch.Tensor:
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    if not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    return x


def empty_nhwc(n, c, h, w, device, dtype=torch.bfloat16):
    return torch.empty((n, c, h, w), dtype=dtype, device=device, memory_format=torch.channels_last)


def shadow(p: torch.Tensor) -> torch.Tensor:
    """bf16 compute copy of an fp32 master, same physical layout.  The fused optimizer keeps
    ``p._shadow`` current; otherwise it is (re)cast whenever the master's version changes."""
    sh = getattr(p, "_shadow", None)
    if sh is not None and getattr(p, "_grad_sink", None) is not None:
        return sh
    ver = p._version
    if sh is None or getattr(p, "_shadow_version", None) != ver or sh.device != p.device:
        with torch.no_grad():
            if p.dim() == 4:
                sh = p.detach().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            else:
                sh = p.detach().to(torch.bfloat16).contiguous()
        p._shadow, p._shadow_version = sh, ver
    return sh


def _grad_target(p: torch.Tensor, like_channels_last: bool = False) -> Tuple[torch.Tensor, bool]:
    """(buffer to accumulate into, routed) -- routed=True means the arena view ``p.grad``."""
    if getattr(p, "_grad_sink", None) is not None and p.grad is not None:
        return p.grad, True
    if like_channels_last:
        return torch.zeros(p.shape, dtype=torch.float32, device=p.device).contiguous(
            memory_format=torch.channels_last), False
    return torch.zeros(p.shape, dtype=torch.float32, device=p.device), False


def _mark_ready(p: torch.Tensor):
    sink = getattr(p, "_grad_sink", None)
    if sink is not None:
        sink.mark_ready(p._arena_index)


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


def _out_size(h, k, s, p):
    return (h + 2 * p - k) // s + 1


def _round_up(a, b):
    return (a + b - 1) // b * b


_replicas_checked = False


class _StatArena:
    """One persistent zeroed buffer per device that hands out the per-layer statistics accumulators of a training step.

    ``begin_step`` (called by FusedDataParallel.forward) re-zeroes the part used since the last call with ONE memset and
    restarts the bump pointer, replacing two ``torch.zeros`` fill kernels per BatchNorm layer per step.  A slice is only
    handed out once between two resets, so it is all zeros when its producer kernel starts; consumers run right after
    the producer in stream order.  Unarmed (no wrapper) or exhausted -> plain ``torch.zeros``."""
    FLOATS = 8 << 20            # 32 MiB: ResNet-101 / inception need ~3.5 M floats per step

    def __init__(self, device):
        self.buf = torch.zeros(self.FLOATS, dtype=torch.float32, device=device)
        self.pos = 0
        self.high = 0
        self.armed = False

    def begin_step(self):
        self.high = max(self.high, self.pos)
        if self.high:
            self.buf[: self.high].zero_()
        self.pos = 0
        self.armed = True

    def take(self, n: int):
        n = (n + 63) // 64 * 64                      # keep every slice 256-byte aligned
        if not self.armed or self.pos + n > self.FLOATS:
            return None
        out = self.buf[self.pos: self.pos + n]
        self.pos += n
        return out


_stat_arenas = {}


def stat_arena(device) -> _StatArena:
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    a = _stat_arenas.get(key)
    if a is None:
        a = _stat_arenas[key] = _StatArena(torch.device("cuda", key))
    return a


def stat_buffer(C: int, device) -> torch.Tensor:
    """Zeroed [STAT_REPLICAS, 2C] accumulator for the statistics kernels."""
    global _replicas_checked
    if not _replicas_checked:
        lib = require_lib()
        assert lib.picd_stat_replicas() == STAT_REPLICAS, "library / Python STAT_REPLICAS mismatch"
        _replicas_checked = True
    n = STAT_REPLICAS * 2 * C
    a = _stat_arenas.get(torch.device(device).index)
    if a is not None:
        t = a.take(n)
        if t is not None:
            return t[:n]
    return torch.zeros(n, dtype=torch.float32, device=device)


# --------------------------------------------------------------------------- conv core
class _ConvInfo:
    __slots__ = ("x", "w_sh", "N", "H", "W", "Cin", "Cout", "R", "S", "stride", "ph", "pw", "OH", "OW",
                 "packed", "orig_w_shape", "s_pack", "Cp")


def _pack_weight(w: torch.Tensor, s: int, Rp: int, Sp: int, Cp: int) -> torch.Tensor:
    """[Cout, C, R, S] fp32 -> bf16 [Cout, Cp, Rp, Sp] channels_last with channel index (a*s+b)*C + c."""
    co, c, r, ss = w.shape
    wp = F.pad(w.detach().float(), (0, Sp * s - ss, 0, Rp * s - r))
    wp = wp.view(co, c, Rp, s, Sp, s).permute(0, 2, 4, 3, 5, 1).reshape(co, Rp, Sp, s * s * c)
    wp = F.pad(wp, (0, Cp - s * s * c))
    return wp.permute(0, 3, 1, 2).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)


def _unpack_wgrad(dwp: torch.Tensor, w_shape, s: int, Rp: int, Sp: int) -> torch.Tensor:
    co, c, r, ss = w_shape
    g = dwp.permute(0, 2, 3, 1)[..., : s * s * c].reshape(co, Rp, Sp, s, s, c)
    g = g.permute(0, 5, 1, 3, 2, 4).reshape(co, c, Rp * s, Sp * s)
    return g[:, :, :r, :ss]


def _conv_forward(x, w, stride, padding, stats, bias, act) -> Tuple[torch.Tensor, _ConvInfo]:
    sh_, sw_ = _pair(stride)
    ph, pw = _pair(padding)
    if sh_ != sw_:
        raise NotImplementedError("picd conv: anisotropic stride")
    info = _ConvInfo()
    N, Cin, H, W = x.shape
    Cout, _, R, S = w.shape
    OH, OW = _out_size(H, R, sh_, ph), _out_size(W, S, sh_, pw)
    info.orig_w_shape, info.packed = tuple(w.shape), False
    if Cin % 8 != 0 or sh_ > 2:
        # first-layer / patch-embedding path: space-to-depth packing makes K tensor-core sized
        s = sh_
        Rp, Sp = -(-R // s), -(-S // s)
        Cp = _round_up(s * s * Cin, 16)
        Hp, Wp = OH + Rp - 1, OW + Sp - 1
        xp = empty_nhwc(N, Cp, Hp, Wp, x.device)
        src_u8 = x.dtype == torch.uint8
        xs = x if src_u8 else nhwc(x)
        call("picd_pack_s2d", xs, xp, N, H, W, Cin, Hp, Wp, Cp, s, ph, pw, int(src_u8))
        w_sh = _pack_weight(w, s, Rp, Sp, Cp)
        info.packed, info.s_pack = True, s
        x, Cin, H, W, R, S, sh_, ph, pw = xp, Cp, Hp, Wp, Rp, Sp, 1, 0, 0
    else:
        x = nhwc(x)
        w_sh = shadow(w)
    if Cout % 8 != 0:
        raise NotImplementedError(f"picd conv: Cout={Cout} must be a multiple of 8")
    y = empty_nhwc(N, Cout, OH, OW, x.device)
    call("picd_conv_fprop", x, w_sh, y, stats, bias, ACT[act], N, H, W, Cin, Cout, R, S, sh_, ph, pw, OH, OW)
    info.x, info.w_sh = x, w_sh
    info.N, info.H, info.W, info.Cin, info.Cout, info.R, info.S = N, H, W, Cin, Cout, R, S
    info.stride, info.ph, info.pw, info.OH, info.OW, info.Cp = sh_, ph, pw, OH, OW, Cin
    return y, info


class GradFork:
    """Side channel between the two consumers of one activation (ResNet block input: first conv + skip path).
    The skip path's backward *deposits* its gradient tensor here; the first conv's dgrad then ADDS its result into
    that tensor with a TMA reduce-add store, so autograd's separate gradient-sum kernel disappears."""
    __slots__ = ("partner",)

    def __init__(self):
        self.partner = None


class _ForkFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, n):
        return tuple(x.view_as(x) for _ in range(n))

    @staticmethod
    def backward(ctx, *gs):
        # branches whose dgrad accumulated into the deposited tensor all hand back that same tensor: count it once
        uniq = {}
        for g in gs:
            if g is not None:
                uniq.setdefault(g.data_ptr(), g)
        vals = list(uniq.values())
        if not vals:
            return None, None
        out = vals[0]
        for g in vals[1:]:
            out = out + g
        return out, None


def fork(x):
    """(x1, x2, fork): two autograd-distinct aliases of x plus the side channel that lets their gradients be fused."""
    x1, x2 = _ForkFn.apply(x, 2)
    return x1, x2, GradFork()


def fork_n(x, n: int):
    """n autograd-distinct aliases of x + the side channel (inception blocks: one activation feeds 3-4 branches).  The
    branch created LAST deposits its input gradient (its backward runs first), the others add theirs into it."""
    return list(_ForkFn.apply(x, n)), GradFork()


class BnRecord:
    """What a conv + BatchNorm(+ReLU) layer leaves behind for whoever produces the gradient of its OUTPUT.

    The BatchNorm backward needs two per-channel sums over (dz, y) before it can write dy.  When the layer's output feeds
    exactly one convolution, that convolution's dgrad kernel -- the kernel that writes dz -- can accumulate the sums in
    its epilogue (``picd_conv_dgrad_bnreduce``): it only needs this layer's raw conv output ``y`` and BN coefficients, which
    this record carries from forward to backward.  The producer fills ``sums`` and remembers which tensor it wrote
    (``dz``); the layer's own backward uses the sums only if autograd hands it exactly that tensor (a second consumer
    of the activation would make autograd sum gradients into a new tensor, and the separate reduction kernel runs).

    Residual layers (two consumers, ReLU behind the skip add) keep the separate reduction: a form of the dgrad that loads
    the skip branch's gradient tile, adds, stores and reduces with the forward's ReLU bit mask was built and measured
    SLOWER than TMA reduce-add + bn_bwd_reduce (3.2 vs 2.3 ms per ResNet-50 step over the 16 layers): three staged tiles
    per output pass make the epilogue, not HBM, the limit of these short-K 1x1 dgrads."""
    __slots__ = ("y", "coef", "relu", "eligible", "sums", "dz")

    def __init__(self):
        self.y = self.coef = self.sums = self.dz = None
        self.relu, self.eligible = False, False


_FUSE_BN_REDUCE = os.environ.get("PICD_FUSE_BN_REDUCE", "1") != "0"     # A/B switch


def _fused_reduce_pays(i: _ConvInfo) -> bool:
    """Where the fused reduction beats dgrad + bn_bwd_reduce (benchmarks/fused_reduce_layers.py, ResNet-50 shapes at batch
    256: 2.65 -> 2.34 ms per step over the 32 eligible layers).  It loses where the dgrad is already bound by its epilogue:
    stride-2 dgrads (four parity launches, each paying the extra tile traffic: 0.119 -> 0.129 ms) and 3x3 layers with
    64-wide tiles (0.120 -> 0.128 ms)."""
    if os.environ.get("PICD_FUSE_BN_REDUCE") == "all":
        return True
    if i.stride != 1:
        return False
    if i.R * i.S > 1 and i.Cin <= 64:
        return False
    return True


def _conv_backward(info: _ConvInfo, dy: torch.Tensor, w: torch.Tensor, need_dx: bool, need_dw: bool, accumulate_into=None,
                   src_bn: Optional["BnRecord"] = None, dbias: Optional[torch.Tensor] = None):
    """Returns (dx or None, dw tensor or None).  dw is None when it was routed into the arena.  ``accumulate_into``:
    an existing bf16 NHWC tensor of dx's shape that dx is added to (and which is returned as dx).  ``src_bn``: record of
    the conv+BN layer that produced this conv's input -- its BatchNorm-backward sums are then reduced by the dgrad."""
    i = info
    dx = None
    if need_dx:
        if i.packed:
            raise NotImplementedError("picd conv: input gradient of a packed first-layer conv")
        if accumulate_into is not None and tuple(accumulate_into.shape) == (i.N, i.Cin, i.H, i.W) and \
                accumulate_into.dtype == torch.bfloat16 and accumulate_into.is_contiguous(memory_format=torch.channels_last):
            dx = accumulate_into
            call("picd_conv_dgrad_accumulate", dy, i.w_sh, dx, i.N, i.H, i.W, i.Cin, i.Cout, i.R, i.S, i.stride, i.ph,
                 i.pw, i.OH, i.OW)
        elif _FUSE_BN_REDUCE and src_bn is not None and src_bn.eligible and src_bn.y is not None and \
                tuple(src_bn.y.shape) == (i.N, i.Cin, i.H, i.W) and _fused_reduce_pays(i):
            dx = empty_nhwc(i.N, i.Cin, i.H, i.W, dy.device)
            coef = src_bn.coef
            sums = stat_buffer(i.Cin, dy.device)
            call("picd_conv_dgrad_bnreduce", dy, i.w_sh, dx, i.N, i.H, i.W, i.Cin, i.Cout, i.R, i.S, i.stride, i.ph, i.pw,
                 i.OH, i.OW, src_bn.y, coef[0], coef[1], coef[2], coef[3], int(src_bn.relu), sums)
            src_bn.sums, src_bn.dz = sums, dx
        else:
            dx = empty_nhwc(i.N, i.Cin, i.H, i.W, dy.device)
            call("picd_conv_dgrad", dy, i.w_sh, dx, i.N, i.H, i.W, i.Cin, i.Cout, i.R, i.S, i.stride, i.ph, i.pw,
                 i.OH, i.OW)
    dw_ret = None
    if need_dw:
        if i.packed:
            dwp = torch.zeros((i.Cout, i.Cp, i.R, i.S), dtype=torch.float32, device=dy.device).contiguous(
                memory_format=torch.channels_last)
            if dbias is not None:
                call("picd_conv_wgrad_bias", i.x, dy, dwp, dbias, i.N, i.H, i.W, i.Cin, i.Cout, i.R, i.S, i.stride, i.ph, i.pw,
                     i.OH, i.OW)
            else:
                call("picd_conv_wgrad", i.x, dy, dwp, i.N, i.H, i.W, i.Cin, i.Cout, i.R, i.S, i.stride, i.ph, i.pw,
                     i.OH, i.OW)
            g = _unpack_wgrad(dwp, i.orig_w_shape, i.s_pack, i.R, i.S)
            tgt, routed = _grad_target(w, True)
            if routed:
                tgt.add_(g)
            else:
                dw_ret = g.contiguous(memory_format=torch.channels_last)
        else:
            tgt, routed = _grad_target(w, True)
            if dbias is not None:
                call("picd_conv_wgrad_bias", i.x, dy, tgt, dbias, i.N, i.H, i.W, i.Cin, i.Cout, i.R, i.S, i.stride, i.ph, i.pw,
                     i.OH, i.OW)
            else:
                call("picd_conv_wgrad", i.x, dy, tgt, i.N, i.H, i.W, i.Cin, i.Cout, i.R, i.S, i.stride, i.ph, i.pw,
                     i.OH, i.OW)
            dw_ret = None if routed else tgt
        if dw_ret is None:
            _mark_ready(w)
    return dx, dw_ret


# --------------------------------------------------------------------------- SyncBN plumbing
def _sync_world(sync_group) -> int:
    if sync_group is False:
        return 1
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return 1
    return dist.get_world_size(None if sync_group is True else sync_group)


def _bn_forward_coeffs(sums, count, gamma, beta, rm, rv, momentum, eps, sync_group):
    C = gamma.numel()
    dev = gamma.device
    coef = torch.empty(4, C, dtype=torch.float32, device=dev)       # scale, shift, mean, invstd
    total = float(count)
    if _sync_world(sync_group) > 1:
        from ..parallel import native_comm
        comm = native_comm.current_comm()
        if comm is not None:
            tot, tot_dev = comm.bn_finalize_sync(sums, count, gamma, beta, rm, rv, momentum, eps, coef)
            return coef, (tot, tot_dev)
        # no peer-mapped slab (e.g. model not wrapped yet): NCCL fallback for the statistics only
        import torch.distributed as dist
        packed = torch.cat([sums.view(-1, 2 * C).sum(0), sums.new_tensor([float(count)])])
        dist.all_reduce(packed, group=None if sync_group is True else sync_group)
        sums = packed[:-1].contiguous()
        total_t = packed[-1:]
        total = float(total_t.item())
    call("picd_bn_finalize", sums, sums.numel() // (2 * C), total, gamma, beta, rm, rv, float(momentum), float(eps),
         coef[0], coef[1], coef[2], coef[3], C)
    return coef, (total, None)


def _bn_allreduce_bwd(sums2, sync_group, C, key=None):
    """[STAT_REPLICAS, 2C] local sums -> the sums bn_bwd_apply should use (folded + all-reduced under SyncBN)."""
    if _sync_world(sync_group) <= 1:
        return sums2
    from ..parallel import native_comm
    comm = native_comm.current_comm()
    if comm is not None:
        return comm.allreduce_small(sums2, 2 * C, key=key)
    import torch.distributed as dist
    out = sums2.view(-1, 2 * C).sum(0)
    dist.all_reduce(out, group=None if sync_group is True else sync_group)
    return out


class _BNCfg:
    __slots__ = ("stride", "padding", "training", "momentum", "eps", "act", "rm", "rv", "sync_group", "fork", "fork_role",
                 "rec", "src_rec")


# Opt-in: let the apply kernel compute scale / shift from the batch sums itself (picd_bn_finalize_apply).  Measured SLOWER
# (+0.3 ms per ResNet-50 step): the prologue runs in each of the ~1000 blocks of every apply launch, which costs more than
# the one tiny bn_finalize kernel it saves.
_FUSE_FINALIZE = os.environ.get("PICD_BN_FUSE_FINALIZE", "0") == "1"
_USE_RELU_MASK = os.environ.get("PICD_BN_MASK", "1") != "0"      # A/B switch: 0 = backward re-reads z for the ReLU sign


def _bn_apply_fwd(ctx, y, gamma, beta, residual, cfg: _BNCfg, sums, count):
    C = gamma.numel()
    P = y.numel() // C
    fuse_finalize = cfg.training and _sync_world(cfg.sync_group) <= 1 and _FUSE_FINALIZE
    total_dev = None           # SyncBN: the exchanged global element count, in device memory
    if fuse_finalize:          # per-replica BN: the apply kernel turns the batch sums into coefficients itself
        coef = torch.empty(4, C, dtype=torch.float32, device=y.device)
        total = float(count)
    elif cfg.training:
        coef, (total, total_dev) = _bn_forward_coeffs(sums, count, gamma, beta, cfg.rm, cfg.rv, cfg.momentum, cfg.eps,
                                                      cfg.sync_group)
    else:
        coef = torch.empty(4, C, dtype=torch.float32, device=y.device)
        call("picd_bn_eval_coeffs", gamma, beta, cfg.rm, cfg.rv, float(cfg.eps), coef[0], coef[1], C)
        total = float(count)
    z = torch.empty_like(y)
    res = None if residual is None else nhwc(residual)
    # residual + ReLU: the backward needs the sign of the OUTPUT (it cannot be recomputed from y alone).  A bit mask
    # (1/16 of the tensor) saves the backward kernels from streaming z again.
    ctx.relu_mask = None
    if res is not None and cfg.act == "relu" and cfg.training and _USE_RELU_MASK:    # (grad mode is off inside Function.forward)
        ctx.relu_mask = torch.empty(P * (C // 8) + 4096, dtype=torch.uint8, device=y.device)
    if fuse_finalize:
        call("picd_bn_finalize_apply", y, sums, sums.numel() // (2 * C), total, gamma, beta, cfg.rm, cfg.rv, float(cfg.momentum),
             float(cfg.eps), coef[0], coef[1], coef[2], coef[3], res, ACT[cfg.act], z, ctx.relu_mask, P, C)
    else:
        call("picd_bn_apply_mask", y, coef[0], coef[1], res, ACT[cfg.act], z, ctx.relu_mask, P, C)
    ctx.bn_total, ctx.bn_act, ctx.bn_training, ctx.bn_total_dev = total, cfg.act, cfg.training, total_dev
    ctx.bn_sync, ctx.has_res = cfg.sync_group, residual is not None
    rec = getattr(cfg, "rec", None)
    ctx.bn_rec = rec
    if rec is not None:
        # the consumer's dgrad may reduce this layer's backward sums when the ReLU mask can be recomputed from y alone
        rec.y, rec.coef, rec.relu = y, coef, cfg.act == "relu"
        rec.eligible = bool(cfg.training and residual is None and cfg.act in ("relu", "none") and C % 8 == 0)
    return z, coef


def _bn_apply_bwd(ctx, dz, y, z, coef, gamma, beta, need_dres):
    """Returns (dy, dres, dgamma, dbeta); dgamma/dbeta are None when routed to the arena."""
    C = gamma.numel()
    P = y.numel() // C
    act = ACT[ctx.bn_act]
    dz = nhwc(dz)
    zmask = z if (act == 1 and ctx.has_res) else None      # ReLU mask needs z only when a residual was added
    dy = torch.empty_like(y)
    dres = torch.empty_like(y) if need_dres else None
    if not ctx.bn_training:
        call("picd_bn_bwd_eval", dz, y, zmask, coef[0], coef[1], act, dy, dres, P, C)
        return dy, dres, None, None
    rmask = getattr(ctx, "relu_mask", None) if zmask is not None else None
    rec = getattr(ctx, "bn_rec", None)
    if rec is not None and rec.sums is not None and rec.dz is not None and rec.dz.data_ptr() == dz.data_ptr() and \
            rec.dz.shape == dz.shape:
        sums2 = rec.sums            # reduced by the dgrad kernel that wrote dz (picd_conv_dgrad_bnreduce)
    else:
        sums2 = stat_buffer(C, y.device)
        call("picd_bn_bwd_reduce_mask", dz, y, zmask, rmask, coef[0], coef[1], coef[2], coef[3], act, sums2, P, C)
    if rec is not None:
        rec.y = rec.coef = rec.sums = rec.dz = None
    gsums = _bn_allreduce_bwd(sums2, ctx.bn_sync, C, key=gamma)
    gt, g_routed = _grad_target(gamma)
    bt, b_routed = _grad_target(beta)
    call("picd_bn_bwd_apply_mask_cnt", dz, y, zmask, rmask, coef[0], coef[1], coef[2], coef[3], act, gsums,
         gsums.numel() // (2 * C), 1.0 / ctx.bn_total, getattr(ctx, "bn_total_dev", None), sums2, STAT_REPLICAS, gt, bt, dy,
         dres, P, C)
    if g_routed:
        _mark_ready(gamma)
    if b_routed:
        _mark_ready(beta)
    return dy, dres, (None if g_routed else gt), (None if b_routed else bt)


# --------------------------------------------------------------------------- fused conv + BN + act
class _ConvBnAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, gamma, beta, residual, cfg: _BNCfg):
        C = w.shape[0]
        stats = stat_buffer(C, x.device) if cfg.training else None
        y, info = _conv_forward(x, w, cfg.stride, cfg.padding, stats, None, None)
        count = y.numel() // C
        z, coef = _bn_apply_fwd(ctx, y, gamma, beta, residual, cfg, stats, count)
        ctx.info = info
        ctx.save_for_backward(w, gamma, beta, y, z, coef)
        ctx.x_needs_grad = x.requires_grad
        ctx.fork, ctx.fork_role = getattr(cfg, "fork", None), getattr(cfg, "fork_role", None)
        ctx.src_rec = getattr(cfg, "src_rec", None)
        return z

    @staticmethod
    def backward(ctx, dz):
        w, gamma, beta, y, z, coef = ctx.saved_tensors
        need = ctx.needs_input_grad
        fork_, role = ctx.fork, ctx.fork_role
        dy, dres, dgamma, dbeta = _bn_apply_bwd(ctx, dz, y, z, coef, gamma, beta, ctx.has_res and need[4])
        if fork_ is not None and role == "deposit_res" and dres is not None:
            fork_.partner = dres
        acc = fork_.partner if (fork_ is not None and role == "accumulate") else None
        # the fused BN-backward reduction needs this dgrad to produce the FINAL gradient of its input: any role but the one
        # that adds into another branch's deposit ("deposit_res" only concerns the residual input, not x)
        dx, dw = _conv_backward(ctx.info, dy, w, need[0], need[1], accumulate_into=acc,
                                src_bn=ctx.src_rec if (fork_ is None or role == "deposit_res") else None)
        if fork_ is not None:
            if role == "deposit_dx" and dx is not None:
                fork_.partner = dx
            # (the deposited tensor stays registered: with more than two branches several dgrads add into it)
        return dx, dw, dgamma, dbeta, dres, None


def conv_bn_act(x, w, gamma, beta, running_mean, running_var, *, stride, padding, training, momentum, eps,
                act, residual, sync_group, grad_fork=None):
    cfg = _BNCfg()
    cfg.stride, cfg.padding, cfg.training, cfg.momentum, cfg.eps = stride, padding, training, momentum, eps
    cfg.act, cfg.rm, cfg.rv, cfg.sync_group = act or "none", running_mean, running_var, sync_group
    cfg.fork, cfg.fork_role = grad_fork if grad_fork is not None else (None, None)
    # BatchNorm-backward reduction fused into the consumer's dgrad (BnRecord): this layer publishes a record on its output,
    # and picks up the record of the layer that produced its input
    cfg.rec = BnRecord() if (training and torch.is_grad_enabled()) else None
    cfg.src_rec = getattr(x, "_bn_rec", None)
    out = _ConvBnAct.apply(x, w, gamma, beta, residual, cfg)
    if cfg.rec is not None:
        out._bn_rec = cfg.rec
    return out


# --------------------------------------------------------------------------- stem: conv + BN + ReLU + 3x3/s2 max-pool
_FUSE_STEM = os.environ.get("PICD_FUSE_STEM", "1") != "0"               # A/B switch


class _ConvBnReluPool(torch.autograd.Function):
    """ResNet / DenseNet stem (reference: nn/classifier.py:36-37 -> torchvision `conv1, bn1, relu, maxpool`): the normalised
    112x112 activation -- the largest tensor of the network -- is never written.  Forward pools relu(BN(y)) straight from the
    conv output; backward regenerates the pooled gradient per input pixel from (dpool, arg-max bytes) inside both BatchNorm
    backward kernels (csrc/ops/stem.cu)."""

    @staticmethod
    def forward(ctx, x, w, gamma, beta, cfg: _BNCfg):
        C = w.shape[0]
        stats = stat_buffer(C, x.device) if cfg.training else None
        y, info = _conv_forward(x, w, cfg.stride, cfg.padding, stats, None, None)
        count = y.numel() // C
        total_dev = None
        if cfg.training:
            coef, (total, total_dev) = _bn_forward_coeffs(stats, count, gamma, beta, cfg.rm, cfg.rv, cfg.momentum, cfg.eps,
                                                          cfg.sync_group)
        else:
            coef = torch.empty(4, C, dtype=torch.float32, device=y.device)
            call("picd_bn_eval_coeffs", gamma, beta, cfg.rm, cfg.rv, float(cfg.eps), coef[0], coef[1], C)
            total = float(count)
        N, _, H, W = y.shape
        OH, OW = _out_size(H, 3, 2, 1), _out_size(W, 3, 2, 1)
        out = empty_nhwc(N, C, OH, OW, y.device)
        arg = torch.empty((N, OH, OW, C), dtype=torch.uint8, device=y.device)
        call("picd_bn_relu_maxpool_fwd", y, coef[0], coef[1], out, arg, N, H, W, C, OH, OW)
        ctx.info, ctx.geom = info, (N, H, W, C, OH, OW)
        ctx.bn_total, ctx.bn_total_dev, ctx.bn_sync = total, total_dev, cfg.sync_group
        ctx.save_for_backward(w, gamma, beta, y, arg, coef)
        return out

    @staticmethod
    def backward(ctx, dpool):
        w, gamma, beta, y, arg, coef = ctx.saved_tensors
        N, H, W, C, OH, OW = ctx.geom
        need = ctx.needs_input_grad
        dpool = nhwc(dpool)
        sums2 = stat_buffer(C, y.device)
        call("picd_bn_relu_maxpool_bwd_reduce", dpool, arg, y, coef[0], coef[1], coef[2], coef[3], sums2, N, H, W, C, OH, OW)
        gsums = _bn_allreduce_bwd(sums2, ctx.bn_sync, C, key=gamma)
        gt, g_routed = _grad_target(gamma)
        bt, b_routed = _grad_target(beta)
        dy = torch.empty_like(y)
        call("picd_bn_relu_maxpool_bwd_apply", dpool, arg, y, coef[0], coef[1], coef[2], coef[3], gsums,
             gsums.numel() // (2 * C), 1.0 / ctx.bn_total, ctx.bn_total_dev, sums2, STAT_REPLICAS, gt, bt, dy, N, H, W, C, OH, OW)
        if g_routed:
            _mark_ready(gamma)
        if b_routed:
            _mark_ready(beta)
        dx, dw = _conv_backward(ctx.info, dy, w, need[0], need[1])
        return dx, dw, (None if g_routed else gt), (None if b_routed else bt), None


def conv_bn_relu_maxpool(x, w, gamma, beta, running_mean, running_var, *, stride, padding, training, momentum, eps,
                         sync_group):
    """max_pool2d(relu(BN(conv(x))), 3, 2, 1).  One fused chain in training mode and under no_grad; an eval-mode BatchNorm
    that still needs gradients (frozen statistics) takes the two separate ops."""
    C = w.shape[0]
    fusable = _FUSE_STEM and C % 8 == 0 and 256 % (C // 8) == 0 and (training or not torch.is_grad_enabled())
    if not fusable:
        z = conv_bn_act(x, w, gamma, beta, running_mean, running_var, stride=stride, padding=padding, training=training,
                        momentum=momentum, eps=eps, act="relu", residual=None, sync_group=sync_group)
        return pool2d(z, 3, 2, 1, mode="max")
    cfg = _BNCfg()
    cfg.stride, cfg.padding, cfg.training, cfg.momentum, cfg.eps = stride, padding, training, momentum, eps
    cfg.act, cfg.rm, cfg.rv, cfg.sync_group = "relu", running_mean, running_var, sync_group
    cfg.fork = cfg.fork_role = cfg.rec = cfg.src_rec = None
    return _ConvBnReluPool.apply(x, w, gamma, beta, cfg)


class _BnAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, gamma, beta, residual, cfg: _BNCfg):
        y = nhwc(y)
        C = gamma.numel()
        P = y.numel() // C
        sums = None
        if cfg.training:
            sums = stat_buffer(C, y.device)
            call("picd_bn_stats", y, sums, P, C)
        z, coef = _bn_apply_fwd(ctx, y, gamma, beta, residual, cfg, sums, P)
        ctx.save_for_backward(gamma, beta, y, z, coef)
        return z

    @staticmethod
    def backward(ctx, dz):
        gamma, beta, y, z, coef = ctx.saved_tensors
        need = ctx.needs_input_grad
        dy, dres, dgamma, dbeta = _bn_apply_bwd(ctx, dz, y, z, coef, gamma, beta, ctx.has_res and need[3])
        return dy, dgamma, dbeta, dres, None


def bn_act(y, gamma, beta, running_mean, running_var, *, training, momentum, eps, act, residual, sync_group):
    cfg = _BNCfg()
    cfg.stride = cfg.padding = None
    cfg.fork = cfg.fork_role = None
    cfg.rec = cfg.src_rec = None
    cfg.training, cfg.momentum, cfg.eps = training, momentum, eps
    cfg.act, cfg.rm, cfg.rv, cfg.sync_group = act or "none", running_mean, running_var, sync_group
    return _BnAct.apply(y, gamma, beta, residual, cfg)


# --------------------------------------------------------------------------- plain conv (+bias)
class _Conv2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, stride, padding):
        y, info = _conv_forward(x, w, stride, padding, None, bias, None)
        ctx.info = info
        ctx.save_for_backward(w, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        w, bias = ctx.saved_tensors
        need = ctx.needs_input_grad
        dy = nhwc(dy)
        db = None
        btgt, b_routed = (None, False)
        if bias is not None and need[2]:
            btgt, b_routed = _grad_target(bias)
        fused_bias = btgt is not None and need[1]          # the wgrad kernel sums dy for the bias as well
        dx, dw = _conv_backward(ctx.info, dy, w, need[0], need[1], dbias=btgt if fused_bias else None)
        if btgt is not None:
            if not fused_bias:
                C = bias.numel()
                call("picd_colsum", dy, btgt, dy.numel() // C, C)
            if b_routed:
                _mark_ready(bias)
            else:
                db = btgt
        return dx, dw, db, None, None


def conv2d(x, w, bias, stride, padding):
    return _Conv2d.apply(x, w, bias, stride, padding)


# --------------------------------------------------------------------------- depthwise conv
class _Depthwise(torch.autograd.Function):
    """Depthwise KxK convolution (groups == channels).  ``pad`` = (top, left, bottom, right) so TF-"same" layers need no
    explicit padding copy."""

    @staticmethod
    def forward(ctx, x, w, stride, pad):
        x = nhwc(x)
        N, C, H, W = x.shape
        K = w.shape[2]
        pt, pl, pb, pr = pad
        OH, OW = (H + pt + pb - K) // stride + 1, (W + pl + pr - K) // stride + 1
        wt = w.detach().reshape(C, K * K).t().contiguous().float()             # tap-major fp32 [K*K, C]
        y = empty_nhwc(N, C, OH, OW, x.device)
        call("picd_depthwise_fprop", x, wt, y, N, H, W, C, OH, OW, K, stride, pt, pl)
        ctx.geom = (N, H, W, C, OH, OW, K, stride, pt, pl)
        ctx.save_for_backward(x, w, wt)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, wt = ctx.saved_tensors
        N, H, W, C, OH, OW, K, stride, pt, pl = ctx.geom
        dy = nhwc(dy)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = empty_nhwc(N, C, H, W, dy.device)
            call("picd_depthwise_dgrad", dy, wt, dx, N, H, W, C, OH, OW, K, stride, pt, pl)
        dw = None
        if ctx.needs_input_grad[1]:
            dwt = torch.zeros(K * K, C, dtype=torch.float32, device=dy.device)
            call("picd_depthwise_wgrad", x, dy, dwt, N, H, W, C, OH, OW, K, stride, pt, pl)
            g = dwt.t().reshape(C, 1, K, K)
            tgt, routed = _grad_target(w, True)
            if routed:
                tgt.add_(g)
                _mark_ready(w)
            else:
                dw = g.contiguous(memory_format=torch.channels_last)
        return dx, dw, None, None


def depthwise_conv2d(x, w, stride, pad):
    """pad: int, (ph, pw) or (top, left, bottom, right)."""
    if isinstance(pad, int):
        pad = (pad, pad, pad, pad)
    elif len(pad) == 2:
        pad = (pad[0], pad[1], pad[0], pad[1])
    s = stride if isinstance(stride, int) else stride[0]
    return _Depthwise.apply(x, w, s, tuple(pad))


# --------------------------------------------------------------------------- linear
_TC_MIN_WORK = 1 << 24


def _use_tc(M, N, K):
    return N % 8 == 0 and K % 8 == 0 and N >= 64 and K >= 64 and M * N * K >= _TC_MIN_WORK


class _LinearTC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, act):
        K = w.shape[1]
        N = w.shape[0]
        x2 = x.reshape(-1, K)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        x2 = x2.contiguous()
        M = x2.shape[0]
        pre = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
        fused_relu = act == "relu"
        call("picd_gemm_nt", x2, shadow(w), pre, b, ACT["relu"] if fused_relu else 0, M, N, K)
        out = pre
        if act not in (None, "none", "relu"):
            out = torch.empty_like(pre)
            call("picd_act_fwd", pre, None, out, M, N, ACT[act])
        ctx.act, ctx.shape = act, x.shape
        ctx.save_for_backward(x2, w, b, pre)
        return out.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w, b, pre = ctx.saved_tensors
        N, K = w.shape
        M = x2.shape[0]
        dy2 = dy.reshape(M, N)
        if dy2.dtype != torch.bfloat16:
            dy2 = dy2.to(torch.bfloat16)
        dy2 = dy2.contiguous()
        bias_done = False
        if ctx.act not in (None, "none"):
            dpre = torch.empty_like(dy2)
            # relu'(pre) == relu'(relu(pre)), so the saved (fused) output serves as the mask source.  The bias gradient is
            # the column sum of dpre: accumulated by the same kernel instead of a second pass over dpre.
            btgt = None
            if b is not None and ctx.needs_input_grad[2] and N % 8 == 0:
                btgt, b_routed = _grad_target(b)
            call("picd_act_bwd_bias", dy2, pre, None, dpre, btgt, M, N, ACT[ctx.act])
            dy2 = dpre
            bias_done = btgt is not None
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K), dtype=torch.bfloat16, device=dy.device)
            call("picd_gemm_nn", dy2, shadow(w), dx, M, N, K)
            dx = dx.view(ctx.shape)
        dw = db = None
        if ctx.needs_input_grad[1]:
            tgt, routed = _grad_target(w)
            if (not bias_done) and b is not None and ctx.needs_input_grad[2]:
                # bias gradient = column sums of dY: the wgrad kernel computes them with a ones-tile MMA (no extra pass)
                btgt, b_routed = _grad_target(b)
                call("picd_gemm_tn_f32_bias", dy2, x2, tgt, btgt, M, N, K)
                bias_done = True
            else:
                call("picd_gemm_tn_f32", dy2, x2, tgt, M, N, K)
            if routed:
                _mark_ready(w)
            else:
                dw = tgt
        if bias_done:
            if b_routed:
                _mark_ready(b)
            else:
                db = btgt
        elif b is not None and ctx.needs_input_grad[2]:
            tgt, routed = _grad_target(b)
            call("picd_colsum", dy2, tgt, M, N)
            if routed:
                _mark_ready(b)
            else:
                db = tgt
        return dx, dw, db, None


class _LinearSmall(torch.autograd.Function):
    """SIMT path for the classifier head (tiny N/K).  Final (activation-free) layers emit fp32 logits."""

    @staticmethod
    def forward(ctx, x, w, b, act):
        N, K = w.shape
        x2 = x.reshape(-1, K)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        x2 = x2.contiguous()
        M = x2.shape[0]
        out_fp32 = act in (None, "none")
        y = torch.empty((M, N), dtype=torch.float32 if out_fp32 else torch.bfloat16, device=x.device)
        call("picd_small_linear_fwd", x2, w, b, y, M, N, K, ACT[act], int(out_fp32))
        ctx.act, ctx.shape = act, x.shape
        ctx.save_for_backward(x2, w, b, y)
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w, b, y = ctx.saved_tensors
        N, K = w.shape
        M = x2.shape[0]
        dy2 = dy.reshape(M, N).contiguous()
        if dy2.dtype not in (torch.float32, torch.bfloat16):
            dy2 = dy2.float()
        scratch = torch.empty((M, N), dtype=torch.float32, device=dy.device)
        dx = torch.empty((M, K), dtype=torch.bfloat16, device=dy.device) if ctx.needs_input_grad[0] else None
        wt, w_routed = _grad_target(w)
        bt, b_routed = (None, False) if b is None else _grad_target(b)
        call("picd_small_linear_bwd", dy2, y, x2, w, scratch, dx, wt, bt, M, N, K, ACT[ctx.act],
             int(dy2.dtype == torch.float32), int(y.dtype == torch.float32))
        if w_routed:
            _mark_ready(w)
        if b_routed:
            _mark_ready(b)
        return (None if dx is None else dx.view(ctx.shape)), (None if w_routed else wt), \
            (None if (b is None or b_routed) else bt), None


def linear(x, w, b, act):
    N, K = w.shape
    M = x.numel() // K
    if _use_tc(M, N, K):
        return _LinearTC.apply(x, w, b, act)
    if act in (None, "none", "relu"):
        return _LinearSmall.apply(x, w, b, act)
    # swish / sigmoid / gelu on tiny tensors (squeeze-excite vectors): native linear, element-wise torch activation
    return ref.activation(_LinearSmall.apply(x, w, b, None), act).to(torch.bfloat16)


# --------------------------------------------------------------------------- pooling
class _Pool2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p, mode, fork_=None):
        ctx.fork = fork_
        x = nhwc(x)
        N, C, H, W = x.shape
        kh, kw = _pair(k)
        OH, OW = _out_size(H, kh, s, p), _out_size(W, kw, s, p)
        out = empty_nhwc(N, C, OH, OW, x.device)
        arg = torch.empty((N, OH, OW, C), dtype=torch.uint8, device=x.device) if mode == 0 else None
        call("picd_pool_fwd", x, out, arg, N, H, W, C, OH, OW, kh, kw, s, p, mode)
        ctx.geom = (N, H, W, C, OH, OW, kh, kw, s, p, mode)
        ctx.save_for_backward(arg) if arg is not None else ctx.save_for_backward()
        return out

    @staticmethod
    def backward(ctx, dy):
        N, H, W, C, OH, OW, kh, kw, s, p, mode = ctx.geom
        arg = ctx.saved_tensors[0] if mode == 0 else None
        dx = empty_nhwc(N, C, H, W, dy.device)
        call("picd_pool_bwd", nhwc(dy), arg, dx, N, H, W, C, OH, OW, kh, kw, s, p, mode)
        if ctx.fork is not None:
            ctx.fork.partner = dx            # sibling branches' dgrads add into this tensor (see fork_n)
        return dx, None, None, None, None, None


def pool2d(x, k, s, p, mode, grad_fork=None):
    if x.shape[1] % 8 != 0:
        return (ref.max_pool2d if mode == "max" else ref.avg_pool2d)(x, k, s, p)
    return _Pool2d.apply(x, k, s, p, 0 if mode == "max" else 1, grad_fork)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = nhwc(x)
        N, C, H, W = x.shape
        out = torch.empty((N, C), dtype=torch.bfloat16, device=x.device)
        if N * ((C // 8 + 31) // 32) < 2 * _sm_count(x.device) and H * W >= 1024:
            # few samples x channels over many pixels: split the pixels over blocks as well (fp32 scratch + finish kernel)
            call("picd_gap_fwd_split", x, torch.zeros((N, C), dtype=torch.float32, device=x.device), out, N, H * W, C)
        else:
            call("picd_gap_fwd", x, out, N, H * W, C)
        ctx.geom = (N, C, H, W)
        return out

    @staticmethod
    def backward(ctx, dy):
        N, C, H, W = ctx.geom
        dx = empty_nhwc(N, C, H, W, dy.device)
        call("picd_gap_bwd", dy.to(torch.bfloat16).contiguous(), dx, N, H * W, C)
        return dx


def global_avg_pool(x):
    if x.shape[1] % 8 != 0:
        return ref.global_avg_pool(x)
    return _GlobalAvgPool.apply(x)


class _ChannelScale(torch.autograd.Function):
    """out = x * s[n, c] * m[n] (squeeze-excite gate, optional stochastic-depth sample mask); csrc/ops/misc.cu."""

    @staticmethod
    def forward(ctx, x, s, m):
        x = nhwc(x)
        N, C, H, W = x.shape
        sb = s.to(torch.bfloat16).contiguous()
        mf = None if m is None else m.float().contiguous()
        out = empty_nhwc(N, C, H, W, x.device)
        call("picd_channel_scale_fwd", x, sb, mf, out, N, H * W, C)
        ctx.save_for_backward(x, sb, mf)
        ctx.s_dtype = s.dtype
        return out

    @staticmethod
    def backward(ctx, dy):
        x, sb, mf = ctx.saved_tensors
        N, C, H, W = x.shape
        dy = nhwc(dy)
        dx = empty_nhwc(N, C, H, W, x.device)
        ds = torch.zeros((N, C), dtype=torch.float32, device=x.device)
        call("picd_channel_scale_bwd", dy, x, sb, mf, dx, ds, N, H * W, C)
        return dx, ds.to(ctx.s_dtype), None


def channel_scale(x, s, sample_mask=None):
    return _ChannelScale.apply(x, s, sample_mask)


# --------------------------------------------------------------------------- loss
class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, weight):
        lg = logits.float().contiguous()
        B, nc = lg.shape
        loss = torch.empty(1, dtype=torch.float32, device=lg.device)
        dlog = torch.empty_like(lg)
        call("picd_cross_entropy", lg, labels.contiguous(), weight, loss, dlog, B, nc)
        ctx.save_for_backward(dlog)
        ctx.in_dtype = logits.dtype
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        (dlog,) = ctx.saved_tensors
        return (dlog * g).to(ctx.in_dtype), None, None


def cross_entropy(logits, labels, weight):
    return _CrossEntropy.apply(logits, labels, None if weight is None else weight.float().contiguous())


# --------------------------------------------------------------------------- activations / dropout / layernorm
class _Activation(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, act):
        xc = x.contiguous() if not (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last)) else x
        C = xc.shape[1] if xc.dim() == 4 else xc.shape[-1]
        out = torch.empty_like(xc)
        call("picd_act_fwd", xc, None, out, xc.numel() // C, C, ACT[act])
        ctx.act = act
        ctx.save_for_backward(xc)
        return out

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        C = x.shape[1] if x.dim() == 4 else x.shape[-1]
        dyc = dy.contiguous(memory_format=torch.channels_last) if x.dim() == 4 and not x.is_contiguous() else dy.contiguous()
        dx = torch.empty_like(x)
        call("picd_act_bwd", dyc, x, None, dx, x.numel() // C, C, ACT[ctx.act])
        return dx, None


def activation(x, act):
    C = x.shape[1] if x.dim() == 4 else x.shape[-1]
    if x.dtype != torch.bfloat16 or C % 8 != 0:
        return ref.activation(x, act)
    return _Activation.apply(x, act)


_step_word = None     # device uint32 view of the step counter while a CUDA graph of the step is recorded (engine/graphed.py)


def set_step_word(t):
    """Route dropout through the graph-replayable kernel variant that reads the step counter from device memory."""
    global _step_word
    _step_word = t


def _dropout_call(src, dst, p, seed, word):
    if word is not None:
        call("picd_dropout_dyn", src, dst, src.numel(), float(p), seed, word)
    else:
        call("picd_dropout", src, dst, src.numel(), float(p), seed)


class _Dropout(torch.autograd.Function):
    _counter = 0

    @staticmethod
    def forward(ctx, x, p):
        _Dropout._counter += 1
        ctx.seed = (int(torch.initial_seed()) * 2654435761 + _Dropout._counter * 40503) & 0xFFFFFFFF
        ctx.p = p
        xc = x.contiguous()
        out = torch.empty_like(xc)
        ctx.word = _step_word
        _dropout_call(xc, out, p, ctx.seed, ctx.word)
        return out

    @staticmethod
    def backward(ctx, dy):
        dyc = dy.contiguous()
        dx = torch.empty_like(dyc)
        _dropout_call(dyc, dx, ctx.p, ctx.seed, ctx.word)
        return dx, None


def dropout(x, p):
    if x.dtype != torch.bfloat16:
        return ref.dropout(x, p, True)
    return _Dropout.apply(x, p)


class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        D = x.shape[-1]
        x2 = x.reshape(-1, D)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        x2 = x2.contiguous()
        M = x2.shape[0]
        y = torch.empty_like(x2)
        stats = torch.empty(2, M, dtype=torch.float32, device=x.device)
        call("picd_layernorm_fwd", x2, gamma, beta, y, stats[0], stats[1], M, D, float(eps))
        ctx.shape = x.shape
        ctx.save_for_backward(x2, gamma, beta, stats)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, gamma, beta, stats = ctx.saved_tensors
        M, D = x2.shape
        dy2 = dy.reshape(M, D).to(torch.bfloat16).contiguous()
        dx = torch.empty_like(x2)
        gt, g_routed = _grad_target(gamma)
        bt, b_routed = _grad_target(beta)
        call("picd_layernorm_bwd", dy2, x2, gamma, stats[0], stats[1], dx, gt, bt, M, D)
        if g_routed:
            _mark_ready(gamma)
        if b_routed:
            _mark_ready(beta)
        return dx.view(ctx.shape), (None if g_routed else gt), (None if b_routed else bt), None


def layer_norm(x, gamma, beta, eps):
    if x.shape[-1] % 8 != 0 or x.shape[-1] > 1024:
        return ref.layer_norm(x, gamma, beta, eps)
    return _LayerNorm.apply(x, gamma, beta, eps)


class _AddLayerNorm(torch.autograd.Function):
    """(s, y) = (x + delta, LayerNorm(x + delta)) in one kernel; the backward adds the gradient of ``s`` (the residual
    stream) to the LayerNorm backward in one kernel as well and hands the SAME tensor to both inputs."""

    @staticmethod
    def forward(ctx, x, delta, gamma, beta, eps):
        D = x.shape[-1]
        x2 = x.reshape(-1, D)
        d2 = delta.reshape(-1, D)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        if d2.dtype != torch.bfloat16:
            d2 = d2.to(torch.bfloat16)
        x2, d2 = x2.contiguous(), d2.contiguous()
        M = x2.shape[0]
        ssum, y = torch.empty_like(x2), torch.empty_like(x2)
        stats = torch.empty(2, M, dtype=torch.float32, device=x.device)
        call("picd_add_layernorm_fwd", x2, d2, gamma, beta, ssum, y, stats[0], stats[1], M, D, float(eps))
        ctx.shape = x.shape
        ctx.save_for_backward(ssum, gamma, beta, stats)
        return ssum.view(x.shape), y.view(x.shape)

    @staticmethod
    def backward(ctx, ds, dy):
        ssum, gamma, beta, stats = ctx.saved_tensors
        M, D = ssum.shape
        if dy is None:
            dy = torch.zeros_like(ssum)
        dy2 = dy.reshape(M, D).to(torch.bfloat16).contiguous()
        ds2 = None if ds is None else ds.reshape(M, D).to(torch.bfloat16).contiguous()
        dx = torch.empty_like(ssum)
        gt, g_routed = _grad_target(gamma)
        bt, b_routed = _grad_target(beta)
        call("picd_layernorm_bwd_add", dy2, ssum, gamma, stats[0], stats[1], ds2, dx, gt, bt, M, D)
        if g_routed:
            _mark_ready(gamma)
        if b_routed:
            _mark_ready(beta)
        dxv = dx.view(ctx.shape)
        return dxv, dxv, (None if g_routed else gt), (None if b_routed else bt), None


def add_layer_norm(x, delta, gamma, beta, eps):
    return _AddLayerNorm.apply(x, delta, gamma, beta, eps)


class _AttentionPacked(torch.autograd.Function):
    """softmax(q k^T / sqrt(64)) v on the tcgen05 kernels, reading q/k/v out of the packed projection [B*S, 3*H*64]."""

    @staticmethod
    def forward(ctx, qkv, B, S, H):
        qkv = qkv.contiguous()
        o = torch.empty((B * S, H * 64), dtype=torch.bfloat16, device=qkv.device)
        lse = torch.empty((B, H, S), dtype=torch.float32, device=qkv.device)
        call("picd_attention_fwd", qkv, o, lse, B, S, H)
        ctx.dims = (B, S, H)
        ctx.save_for_backward(qkv, o, lse)
        return o

    @staticmethod
    def backward(ctx, do):
        qkv, o, lse = ctx.saved_tensors
        B, S, H = ctx.dims
        do = do.to(torch.bfloat16).contiguous()
        dqkv = torch.empty_like(qkv)
        delta = torch.empty_like(lse)
        call("picd_attention_bwd", qkv, o, do, lse, delta, dqkv, B, S, H)
        return dqkv, None, None, None


def attention_packed_supported(qkv, S: int, H: int) -> bool:
    return qkv.dtype == torch.bfloat16 and qkv.shape[-1] == 3 * H * 64 and 1 <= S <= 256


def attention_packed(qkv, B: int, S: int, H: int):
    return _AttentionPacked.apply(qkv, B, S, H)


def attention(q, k, v, scale):
    # generic [B, heads, S, D] entry (head dim != 64 or S > 256): library kernel.  ViT-B/16 goes through
    # attention_packed (csrc/ops/attention.cu) instead.
    return F.scaled_dot_product_attention(q, k, v, scale=scale)


# --------------------------------------------------------------------------- data
def synthetic_batch(batch, channels, height, width, num_classes, seed, step, device, dtype):
    img = torch.empty((batch, channels, height, width), dtype=dtype, device=device,
                      memory_format=torch.channels_last)
    labels = torch.empty(batch, dtype=torch.int64, device=device)
    base = (seed * 0x9E3779B1 + step * 0x85EBCA77) & 0xFFFFFFFF
    if dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("synthetic_batch: dtype must be fp32 or bf16")
    with torch.cuda.device(device):
        call("picd_synthetic_batch", img, labels, batch, height, width, channels, num_classes, base,
             int(dtype == torch.float32))
    return img, labels


def normalize_u8(images_u8, dtype):
    n, h, w, c = images_u8.shape
    src = images_u8.contiguous()
    out = torch.empty((n, c, h, w), dtype=dtype, device=src.device, memory_format=torch.channels_last)
    call("picd_normalize_u8", src, out, src.numel(), c, int(dtype == torch.float32))
    return out

# --------------------------------------------------------------------------- channel concat (inception blocks, K12)
def _concat_call(parts, wide, split):
    n = len(parts)
    ptrs = (ctypes.c_void_p * n)(*[t.data_ptr() for t in parts])
    chans = (ctypes.c_int * n)(*[t.shape[1] for t in parts])
    P = wide.numel() // wide.shape[1]
    call("picd_concat_nhwc", ctypes.cast(ptrs, ctypes.c_void_p).value, ctypes.cast(chans, ctypes.c_void_p).value, n, wide, P,
         int(split))


class _ConcatNHWC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, *xs):
        xs = [nhwc(x) for x in xs]
        N, _, H, W = xs[0].shape
        ctx.channels = [x.shape[1] for x in xs]
        wide = empty_nhwc(N, sum(ctx.channels), H, W, xs[0].device)
        _concat_call(xs, wide, 0)
        return wide

    @staticmethod
    def backward(ctx, dwide):
        dwide = nhwc(dwide)
        N, _, H, W = dwide.shape
        parts = [empty_nhwc(N, c, H, W, dwide.device) for c in ctx.channels]
        _concat_call(parts, dwide, 1)
        return tuple(parts)


def concat_channels(xs):
    if len(xs) > 8 or any(x.shape[1] % 8 != 0 for x in xs):
        return torch.cat(xs, 1).contiguous(memory_format=torch.channels_last)
    return _ConcatNHWC.apply(*xs)


def top1_hits(logits, labels):
    """int64 [B]: 1 where argmax(logits) == label, one launch for the whole batch."""
    lg = logits.detach()
    if lg.dtype not in (torch.bfloat16, torch.float32):
        lg = lg.float()
    lg = lg.contiguous()
    lab = labels.to(torch.int64).contiguous()
    out = torch.empty(lg.shape[0], dtype=torch.int64, device=lg.device)
    call("picd_top1_hits", lg, int(lg.dtype == torch.bfloat16), lab, out, lg.shape[0], lg.shape[1])
    return out
