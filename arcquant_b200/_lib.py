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
    "arcq_add_rmsnorm_quantize_x": (_I32, [_VP, _VP, _VP, _F32, _I64, _I32, _VP, _I32, _VP, _VP, _VP, _VP]),
    "arcq_gemm_nvfp4": (_I32, [_VP, _VP, _VP, _VP, _I64, _I64, _I64, _F32, _VP, _VP, _VP]),
    "arcq_gemm_nvfp4_ex": (_I32, [_VP, _VP, _VP, _VP, _I64, _I64, _I64, _F32, _VP, _VP, _I32, _VP, _I64, _VP]),
    "arcq_gemm_nvfp4_ex2": (_I32, [_VP, _VP, _VP, _VP, _I64, _I64, _I64, _F32, _VP, _VP, _VP, _I32, _VP, _I64, _VP]),
    "arcq_absmax": (_I32, [_VP, _I64, _VP, _VP]),
    "arcq_quantize_x_scaled": (_I32, [_VP, _I64, _I32, _VP, _I32, _I32, _VP, _I32, _VP, _VP, _VP, _VP]),
    "arcq_set_option": (_I32, [ctypes.c_char_p, ctypes.c_char_p]),
    "arcq_plan_entries": (_I64, [_I32]),
    "arcq_plan_build": (_I32, [_VP, _I32, _VP]),
    "arcq_plan_max_conflict": (_I32, [_VP, _I32, _I32, _VP]),
    "arcq_quantize_x_ex": (_I32, [_VP, _I64, _I32, _VP, _I32, _I32, _VP, _VP, _I32, _VP, _VP, _VP, _VP]),
    "arcq_linear_noalloc_bytes": (_I64, [_I64, _I32, _I32]),
    "arcq_linear_noalloc": (_I32, [_VP, _I64, _I32, _VP, _I32, _I32, _VP, _VP, _I64, _F32, _VP, _VP, _VP]),
    "arcq_peer_alloc": (_I32, [_I64, _VP]),
    "arcq_peer_free": (_I32, [_VP]),
    "arcq_peer_export": (_I32, [_VP, _VP]),
    "arcq_peer_open": (_I32, [_VP, _VP]),
    "arcq_peer_close": (_I32, [_VP]),
    "arcq_peer_enable": (_I32, [_I32]),
    "arcq_peer_status": (_I32, [_VP, _VP]),
    "arcq_gemm_nvfp4_scatter": (_I32, [_VP, _VP, _VP, _VP, _I64, _I64, _I64, _F32, _VP, _VP, _VP, _I32, _I32, _I64, _I64, _VP]),
    "arcq_peer_reduce": (_I32, [_VP, _VP, _I32, _I32, _I32, _I64, _I64, _I64, _VP, _I64, _VP]),
    "arcq_peer_allgather": (_I32, [_VP, _I64, _VP, _I64, _VP, _I64, _VP, _I64, _VP, _I32, _I32, _I32, _VP]),
    "arcq_launch_count": (_I64, []),
    "arcq_last_error": (ctypes.c_char_p, []),
    "arcq_version": (ctypes.c_char_p, []),
}


def lib_path() -> str:
    return _build.LIB


def _ensure_built() -> str:
    """Path of the library to load: the in-tree build, rebuilt first when it is absent or STALE (sources changed since
    the stamp) and a compiler is present.  On the GPU box there is no reason to rebuild: the snapshot's .so is current."""
    path = os.environ.get("ARCQ_LIB")  # experiment builds (scripts/build_variant.py); loader-level, never read at launch
    if path:
        return path
    if (not _build.is_current() and _build.can_build()) or not os.path.exists(_build.LIB):
        _build.build()
    return _build.LIB


def load() -> ctypes.CDLL:
    """ctypes handle of libarcq_b200.so (building first if the in-tree .so is absent or stale and nvcc is available)."""
    global _L
    if _L is not None:
        return _L
    path = _ensure_built()
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


_EXT = False  # False = not tried yet, None = unavailable


def load_ext():
    """The compiled pybind11/torch module over the same C ABI (csrc/agemm_ext.cpp), or None when it cannot be imported
    (then agemm.py binds the C ABI through ctypes instead -- same kernels, ~15 us more host time per call)."""
    global _EXT
    if _EXT is not False:
        return _EXT
    _EXT = None
    if os.environ.get("ARCQ_LIB") or os.environ.get("ARCQ_NO_EXT") == "1":
        return None
    try:
        _ensure_built()
        import importlib.util
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        load()  # libarcq_b200.so by path first, so the extension's DT_NEEDED entry resolves to the same handle
        spec = importlib.util.spec_from_file_location(_build.EXT_NAME, _build.EXT)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _EXT = mod
    except Exception as e:  # pragma: no cover - depends on the installed torch
        import warnings
        warnings.warn("arcquant_b200: compiled agemm module unavailable (%s); using the ctypes binding" % (e,))
        _EXT = None
    return _EXT


def last_error() -> str:
    return load().arcq_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError("%s failed (status %d): %s" % (what, rc, last_error()))
