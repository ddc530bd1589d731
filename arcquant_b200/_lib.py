"""ctypes binding of libarcq_b200.so -- the C ABI declared in include/arcq.h.

The product path has no CPU fallback: if the library is missing it is built in-tree (nvcc), and if
that fails, or no CUDA device is present when a compute entry point is called, the call raises.
"""
from __future__ import annotations

import ctypes
import os

from . import build as _build

_L = None

_VP, _I64, _I32, _F32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float

# name -> (restype, argtypes); must list every symbol include/arcq.h declares
SIGNATURES = {
    "arcq_flavour_for_kq": (_I32, [_I32]),
    "arcq_sf_bytes": (_I64, [_I64, _I64]),
    "arcq_sf_offset": (_I64, [_I64, _I64, _I64]),
    "arcq_quantize_x": (_I32, [_VP, _I64, _I32, _VP, _I32, _I32, _VP, _VP, _VP]),
    "arcq_quantize_w": (_I32, [_VP, _I64, _I32, _VP, _I32, _I32, _VP, _VP, _VP]),
    "arcq_rmsnorm_quantize_x": (_I32, [_VP, _VP, _F32, _I64, _I32, _VP, _I32, _VP, _VP, _VP]),
    "arcq_gemm_nvfp4": (_I32, [_VP, _VP, _VP, _VP, _I64, _I64, _I64, _F32, _VP, _VP, _VP]),
    "arcq_launch_count": (_I64, []),
    "arcq_last_error": (ctypes.c_char_p, []),
    "arcq_version": (ctypes.c_char_p, []),
}


def lib_path() -> str:
    return _build.LIB


def load() -> ctypes.CDLL:
    """Load (building first if the in-tree .so is absent or stale and nvcc is available)."""
    global _L
    if _L is not None:
        return _L
    path = os.environ.get("ARCQ_LIB") or _build.LIB  # ARCQ_LIB: experiment builds (scripts/build_variant.py)
    if not os.path.exists(path) or (os.environ.get("ARCQ_REBUILD") == "1"):
        path = _build.build()
    try:
        L = ctypes.CDLL(path)
    except OSError as e:  # fail loudly: there is no fallback implementation
        raise ImportError("arcquant_b200: cannot load %s (%s); run `python -m arcquant_b200.build`" % (path, e))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _L = L
    return L


def last_error() -> str:
    return load().arcq_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError("%s failed (status %d): %s" % (what, rc, last_error()))
