"""Host-side mirror of the reference's pybind11 module ``agemm`` (kernels/src/bindings.cpp:547-582).

Same function names, argument names, shapes, dtypes and error behaviour as the reference for the four
hot entry points; every call is one launch of a hand-written sm_100a kernel through the C ABI of
``include/arcq.h`` on the CURRENT torch stream.  Differences from the reference, all deliberate:

* kernels run on ``torch.cuda.current_stream()`` (the reference uses the legacy default stream);
* arguments are validated (device, dtype, contiguity, shapes) -> ``RuntimeError``;
* ``KQ`` is generic (any multiple of 16 with ``(KQ+KE) % 64 == 0``), not the reference's table
  (bindings.cpp:141-160); for KQ in the table the reference's flavour (16- vs 32-wide ARC interleave
  and residual formula) is reproduced bit-exactly;
* ``matmul`` also accepts a 0-dim CUDA tensor as ``scale`` and then never synchronises the host.

There is no CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import torch

from . import _lib

__all__ = [
    "matmul", "reorder_quantize_x", "reorder_quantize_w", "rmsnorm_quantize_x", "sf_bytes",
    "flavour_for_kq", "launch_count",
]


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _need(t, name, dtype, ndim=None):
    if not isinstance(t, torch.Tensor):
        raise RuntimeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (arcquant_b200 has no CPU path)" % name)
    if t.dtype != dtype:
        raise RuntimeError("%s must have dtype %s, got %s" % (name, dtype, t.dtype))
    if ndim is not None and t.dim() != ndim:
        raise RuntimeError("%s must be %d-dimensional" % (name, ndim))
    if not t.is_contiguous():
        raise RuntimeError("%s must be contiguous" % name)


def sf_bytes(rows: int, K: int) -> int:
    """Scale-factor allocation of the reference (bindings.cpp:83-95): (rows//128 + 1)*128*K//16."""
    return int(_lib.load().arcq_sf_bytes(rows, K))


def flavour_for_kq(KQ: int) -> int:
    return int(_lib.load().arcq_flavour_for_kq(KQ))


def launch_count() -> int:
    return int(_lib.load().arcq_launch_count())


class _on:
    """`with torch.cuda.device(dev)` only when dev is not already current (the context manager costs microseconds)."""
    __slots__ = ("ctx",)

    def __init__(self, dev):
        self.ctx = None if dev.index == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx is not None:
            self.ctx.__exit__(*a)


def _sf_bytes(rows: int, K: int) -> int:  # == arcq_sf_bytes, without the ctypes round trip
    return (rows // 128 + 1) * 128 * K // 16


def _quant(fn_name, X, reorder_index, KE, flavour):
    L = _lib.load()
    _need(X, "X", torch.bfloat16, 2)
    _need(reorder_index, "reorder_index", torch.int16, 1)
    M, KQ = X.shape
    if reorder_index.numel() != KQ or reorder_index.device != X.device:
        raise RuntimeError("reorder_index must hold %d int16 entries on %s" % (KQ, X.device))
    KE = int(KE)
    K = KQ + KE
    with _on(X.device):
        q = torch.empty((M, K // 2), dtype=torch.uint8, device=X.device)
        sf = torch.empty((_sf_bytes(M, K),), dtype=torch.uint8, device=X.device)
        rc = getattr(L, fn_name)(X.data_ptr(), M, KQ, reorder_index.data_ptr(), KE, flavour,
                                 q.data_ptr(), sf.data_ptr(), _stream(X))
    if rc:
        _lib.check(rc, fn_name)
    return q, sf


def reorder_quantize_x(X, reorder_index, KE, flavour: int = -1):
    """Reorder and quantize activation, return K + KE channels (bindings.cpp:122-163)."""
    return _quant("arcq_quantize_x", X, reorder_index, KE, flavour)


def reorder_quantize_w(W, reorder_index, KE, flavour: int = -1):
    """Reorder and quantize weight, return K + KE channels (bindings.cpp:170-210)."""
    return _quant("arcq_quantize_w", W, reorder_index, KE, flavour)


def rmsnorm_quantize_x(X, W, eps, reorder_index, KE):
    """Rmsnorm and quantize activation, return K + KE channels (bindings.cpp:216-254)."""
    L = _lib.load()
    _need(X, "X", torch.bfloat16, 2)
    _need(W, "W", torch.bfloat16, 1)
    _need(reorder_index, "reorder_index", torch.int16, 1)
    M, H = X.shape
    if W.numel() != H or reorder_index.numel() != H:
        raise RuntimeError("W and reorder_index must have %d entries" % H)
    KE = int(KE)
    K = H + KE
    with _on(X.device):
        q = torch.empty((M, K // 2), dtype=torch.uint8, device=X.device)
        sf = torch.empty((_sf_bytes(M, K),), dtype=torch.uint8, device=X.device)
        rc = L.arcq_rmsnorm_quantize_x(X.data_ptr(), W.data_ptr(), float(eps), M, H, reorder_index.data_ptr(),
                                       KE, q.data_ptr(), sf.data_ptr(), _stream(X))
    _lib.check(rc, "arcq_rmsnorm_quantize_x")
    return q, sf


def matmul(A, B, SFA, SFB, scale, out=None):
    """nvfp4 matmul, return bfloat16 tensor (bindings.cpp:99-120): scale * (A.SFA) @ (B.SFB)^T."""
    L = _lib.load()
    _need(A, "A", torch.uint8, 2)
    _need(B, "B", torch.uint8, 2)
    _need(SFA, "SFA", torch.uint8)
    _need(SFB, "SFB", torch.uint8)
    M, K2 = A.shape
    N = B.shape[0]
    if B.shape[1] != K2:
        raise RuntimeError("A and B must share the packed K dimension (%d vs %d)" % (K2, B.shape[1]))
    K = 2 * K2
    need_a = ((M + 127) // 128) * 128 * K // 16
    need_b = ((N + 127) // 128) * 128 * K // 16
    if SFA.numel() < need_a or SFB.numel() < need_b:
        raise RuntimeError("scale-factor buffers too small for M=%d N=%d K=%d" % (M, N, K))
    alpha_dev = 0
    alpha = 0.0
    if isinstance(scale, torch.Tensor) and scale.is_cuda:
        if scale.dtype != torch.float32 or scale.numel() != 1:
            raise RuntimeError("a device `scale` must be one float32")
        alpha_dev = scale.data_ptr()
    else:
        alpha = float(scale)
    with _on(A.device):
        if out is None:
            out = torch.empty((M, N), dtype=torch.bfloat16, device=A.device)
        else:
            _need(out, "out", torch.bfloat16, 2)
        rc = L.arcq_gemm_nvfp4(A.data_ptr(), B.data_ptr(), SFA.data_ptr(), SFB.data_ptr(), M, N, K, alpha,
                               alpha_dev, out.data_ptr(), _stream(A))
    _lib.check(rc, "arcq_gemm_nvfp4")
    return out


def _not_on_this_path(name):
    def f(*a, **k):
        raise NotImplementedError(
            "agemm.%s is the reference's int4 KV-cache path (bindings.cpp:258-543), outside the NVFP4+ARC hot path" % name)
    f.__name__ = name
    return f


for _n in ("batch_decode_i4", "init_kv_i4", "append_kv_i4", "batch_decode_f16", "init_kv_f16", "append_kv_f16"):
    globals()[_n] = _not_on_this_path(_n)
