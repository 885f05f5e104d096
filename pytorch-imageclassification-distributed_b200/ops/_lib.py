"""ctypes loader for ``libpicd_b200.so`` (built in-tree by ``build.py``).

The kernel library exposes a plain C ABI (``include/picd_b200.h``): every entry
point takes raw device pointers, sizes and a ``cudaStream_t`` and returns an
``int`` status (0 = ok).  Binding through ctypes keeps the build a pure ``nvcc``
invocation (seconds, no torch headers) and lets non-Python hosts link the same
library.
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB_NAME = "libpicd_b200.so"
_lock = threading.Lock()
_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[str] = None


def lib_path() -> str:
    return os.environ.get("PICD_B200_LIB", os.path.join(_PKG_DIR, "lib", _LIB_NAME))


def _try_load() -> None:
    global _lib, _load_error
    with _lock:
        if _lib is not None or _load_error is not None:
            return
        path = lib_path()
        if not os.path.exists(path):
            _load_error = f"{path} does not exist (run `python build.py` or __graft_entry__.build())"
            return
        try:
            _lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
            _lib.picd_last_error.restype = ctypes.c_char_p
            _lib.picd_version.restype = ctypes.c_int
        except OSError as e:  # pragma: no cover - depends on the box
            _load_error = f"dlopen({path}) failed: {e}"


def lib_available() -> bool:
    _try_load()
    return _lib is not None


def require_lib() -> ctypes.CDLL:
    """Return the loaded library or raise -- CUDA tensors never fall back silently."""
    _try_load()
    if _lib is None:
        raise RuntimeError(
            "picd_b200: the sm_100a kernel library is required for CUDA tensors but is not "
            f"loadable: {_load_error}")
    return _lib


class PicdError(RuntimeError):
    pass


def check(status: int, what: str) -> None:
    if status != 0:
        lib = require_lib()
        msg = lib.picd_last_error()
        raise PicdError(f"{what} failed (status {status}): {msg.decode() if msg else '?'}")
