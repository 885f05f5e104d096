"""The allocation-free classifier-head backward, ``picd_small_linear_bwd_noalloc``.

Some hosts of this library are not allowed to create device memory at the moment they need the head's
backward: a process that gave its whole memory budget to its own pool at start-up and bans ``cudaMalloc``
afterwards, a plug-in that lives under another framework's allocator, a loop that cannot pay for
``cudaMalloc``'s device-wide synchronisation.  They can supply inputs and outputs but not the ``dpre`` scratch
that ``picd_small_linear_bwd`` wants, so this variant keeps that scratch in a 1 MiB array inside the library.

The test drives both entry points through the plain C ABI (ctypes, raw pointers) and requires the same
``dx`` / ``dW`` / ``db`` within the tolerance the other small-linear tests use (they are in fact the same
kernels, so the results should be identical).  It does so twice in a row with different shapes and data: the
built-in array is shared by all calls and still holds the first call's values when the second one starts, and
the second result must not depend on them.  It also checks that a problem larger than the built-in array is
refused with a readable message instead of being launched.

This file sorts last on purpose, so a run that stops at the first failure has already run everything else.
"""
import ctypes

import pytest

pytestmark = pytest.mark.gpu

TOL = 3e-2          # same bound as the small_linear cases in benchmarks/gpu_selftest.py


def _lib():
    from picd_b200.ops import require_lib
    lib = require_lib()
    vp, i = ctypes.c_void_p, ctypes.c_int
    lib.picd_small_linear_bwd.argtypes = [vp] * 8 + [i] * 6 + [vp]
    lib.picd_small_linear_bwd.restype = i
    lib.picd_small_linear_bwd_noalloc.argtypes = [vp] * 7 + [i] * 6 + [vp]
    lib.picd_small_linear_bwd_noalloc.restype = i
    lib.picd_small_linear_noalloc_max_floats.restype = i
    lib.picd_last_error.restype = ctypes.c_char_p
    return lib


def _relerr(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6))


def _one_case(lib, M, K, N, act, seed):
    import torch
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(M, K, generator=g).to(dev).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(dev)
    y = torch.randn(M, N, generator=g).to(dev).to(torch.bfloat16)       # forward output (ReLU mask source)
    dy = torch.randn(M, N, generator=g).to(dev).to(torch.bfloat16)
    stream = torch.cuda.current_stream().cuda_stream

    # existing entry point: the caller brings the scratch
    scratch = torch.empty(M, N, dtype=torch.float32, device=dev)
    dx_a = torch.empty(M, K, dtype=torch.bfloat16, device=dev)
    dw_a = torch.zeros(N, K, dtype=torch.float32, device=dev)
    db_a = torch.zeros(N, dtype=torch.float32, device=dev)
    rc = lib.picd_small_linear_bwd(dy.data_ptr(), y.data_ptr(), x.data_ptr(), w.data_ptr(), scratch.data_ptr(),
                                   dx_a.data_ptr(), dw_a.data_ptr(), db_a.data_ptr(), M, N, K, act, 0, 0, stream)
    assert rc == 0, lib.picd_last_error()

    # new entry point: inputs and outputs only
    dx_b = torch.empty(M, K, dtype=torch.bfloat16, device=dev)
    dw_b = torch.zeros(N, K, dtype=torch.float32, device=dev)
    db_b = torch.zeros(N, dtype=torch.float32, device=dev)
    rc = lib.picd_small_linear_bwd_noalloc(dy.data_ptr(), y.data_ptr(), x.data_ptr(), w.data_ptr(), dx_b.data_ptr(),
                                           dw_b.data_ptr(), db_b.data_ptr(), M, N, K, act, 0, 0, stream)
    assert rc == 0, lib.picd_last_error()
    torch.cuda.synchronize()

    for name, a, b in (("dx", dx_a, dx_b), ("dw", dw_a, dw_b), ("db", db_a, db_b)):
        assert torch.isfinite(b.float()).all(), name
        assert _relerr(b, a) <= TOL, f"{name}: {_relerr(b, a)}"
    # and both agree with plain PyTorch in fp32
    dpre = dy.float() * ((y.float() > 0).float() if act == 1 else 1.0)
    assert _relerr(dx_b, dpre @ w) <= TOL
    assert _relerr(dw_b, dpre.t() @ x.float()) <= TOL
    assert _relerr(db_b, dpre.sum(0)) <= TOL


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    lib = _lib()
    _one_case(lib, M=96, K=2048, N=128, act=1, seed=1)     # leaves 96*128 floats behind in the shared array
    _one_case(lib, M=37, K=64, N=32, act=0, seed=2)        # smaller problem, must not see them


def test_noalloc_rejects_oversize_problems():
    import torch
    lib = _lib()
    cap = lib.picd_small_linear_noalloc_max_floats()
    assert cap == 1 << 18
    M, N, K = 1024, cap // 1024 + 1, 8                    # M*N just above the built-in capacity
    dev = torch.device("cuda", 0)
    dummy = torch.zeros(16, device=dev)
    rc = lib.picd_small_linear_bwd_noalloc(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), None,
                                           dummy.data_ptr(), dummy.data_ptr(), M, N, K, 0, 0, 0,
                                           torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.picd_last_error().decode()
    assert "exceed" in msg and str(cap) in msg
    torch.cuda.synchronize()                               # nothing was launched, the context is healthy
