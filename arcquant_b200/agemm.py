"""Host-side mirror of the reference's pybind11 module ``agemm`` (kernels/src/bindings.cpp:547-582).

Same function names, argument names, shapes, dtypes and error behaviour as the reference for the four hot entry
points; every call is one launch of a hand-written sm_100a kernel through the C ABI of ``include/arcq.h`` on the
CURRENT torch stream.  The functions exported here ARE the functions of the compiled pybind11/torch module
``_agemm_ext`` (csrc/agemm_ext.cpp -- what INTEGRATION.md §B describes); only when that module cannot be imported
are they the ctypes bindings below (same C ABI, same kernels, more host time per call).  Differences from the
reference, all deliberate:

* kernels run on ``torch.cuda.current_stream()`` (the reference uses the legacy default stream);
* arguments are validated (device, dtype, contiguity, shapes) -> ``RuntimeError``;
* ``KQ`` is generic (any multiple of 16 with ``(KQ+KE) % 64 == 0``), not the reference's table
  (bindings.cpp:141-160); for KQ in the table the reference's flavour (16- vs 32-wide ARC interleave
  and residual formula) is reproduced bit-exactly.  NOTE the reference's own table mismatch: ``rmsnorm_quantize_x``
  is ALWAYS the 16-wide flavour (bindings.cpp:243) while ``reorder_quantize_w`` picks the 32-wide one for
  KQ in {3584, 18944, 27648, 28672} (:199) -- weights consumed after ``rmsnorm_quantize_x`` must be quantised with
  ``flavour=FLAVOUR_16``;
* ``matmul`` also accepts a 1-element CUDA tensor as ``scale`` and then never synchronises the host;
* extras: ``matmul(..., col_scale=, epilogue=, out=)`` (several weights in one launch, SwiGLU
  epilogue), ``reorder_quantize_x_scaled`` / ``absmax`` (the callers' per-tensor pre-scale fused in), ``set_option``.

There is no CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import torch

from . import _lib

__all__ = [
    "matmul", "reorder_quantize_x", "reorder_quantize_w", "rmsnorm_quantize_x", "add_rmsnorm_quantize_x", "reorder_quantize_x_scaled", "absmax",
    "sf_bytes", "flavour_for_kq", "launch_count", "set_option", "use_plans", "fuse_weights", "interleave_gate_up",
    "EPILOGUE_NONE", "EPILOGUE_SWIGLU", "PRESCALE_FP32", "PRESCALE_BF16_DIVISOR", "FLAVOUR_16", "FLAVOUR_32", "BACKEND",
]

EPILOGUE_NONE, EPILOGUE_SWIGLU = 0, 1
PRESCALE_FP32, PRESCALE_BF16_DIVISOR = 0, 1
FLAVOUR_16, FLAVOUR_32 = 0, 1


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


def _sf_bytes(rows: int, K: int) -> int:  # == arcq_sf_bytes, without the ctypes round trip
    return (rows // 128 + 1) * 128 * K // 16


# ---- ctypes binding of the C ABI (fallback when the compiled module is unavailable) ------------------------------

def _c_sf_bytes(rows: int, K: int) -> int:
    """Scale-factor allocation of the reference (bindings.cpp:83-95): (rows//128 + 1)*128*K//16."""
    return int(_lib.load().arcq_sf_bytes(rows, K))


def _c_flavour_for_kq(KQ: int) -> int:
    return int(_lib.load().arcq_flavour_for_kq(KQ))


def _c_launch_count() -> int:
    return int(_lib.load().arcq_launch_count())


def _c_set_option(name: str, value: str = "") -> None:
    _lib.check(_lib.load().arcq_set_option(name.encode(), str(value).encode()), "arcq_set_option")


def _quant(fn_name, X, reorder_index, KE, flavour, amax=None, mode=0):
    L = _lib.load()
    _need(X, "X", torch.bfloat16, 2)
    _need(reorder_index, "reorder_index", torch.int16, 1)
    M, KQ = X.shape
    if reorder_index.numel() != KQ or reorder_index.device != X.device:
        raise RuntimeError("reorder_index must hold %d int16 entries on %s" % (KQ, X.device))
    KE = int(KE)
    K = KQ + KE
    with torch.cuda.device(X.device):
        q = torch.empty((M, K // 2), dtype=torch.uint8, device=X.device)
        sf = torch.empty((_sf_bytes(M, K),), dtype=torch.uint8, device=X.device)
        if fn_name == "arcq_quantize_x_scaled":
            scale = torch.empty((1,), dtype=torch.float32, device=X.device)
            rc = L.arcq_quantize_x_scaled(X.data_ptr(), M, KQ, reorder_index.data_ptr(), KE, flavour, amax.data_ptr(),
                                          int(mode), q.data_ptr(), sf.data_ptr(), scale.data_ptr(), _stream(X))
        else:
            rc = getattr(L, fn_name)(X.data_ptr(), M, KQ, reorder_index.data_ptr(), KE, flavour,
                                     q.data_ptr(), sf.data_ptr(), _stream(X))
    if rc:
        _lib.check(rc, fn_name)
    return (q, sf, scale) if fn_name == "arcq_quantize_x_scaled" else (q, sf)


def _c_reorder_quantize_x(X, reorder_index, KE, flavour: int = -1):
    """Reorder and quantize activation, return K + KE channels (bindings.cpp:122-163)."""
    return _quant("arcq_quantize_x", X, reorder_index, KE, flavour)


def _c_reorder_quantize_w(W, reorder_index, KE, flavour: int = -1):
    """Reorder and quantize weight, return K + KE channels (bindings.cpp:170-210)."""
    return _quant("arcq_quantize_w", W, reorder_index, KE, flavour)


def _c_absmax(X):
    _need(X, "X", torch.bfloat16)
    with torch.cuda.device(X.device):
        out = torch.empty((1,), dtype=torch.float32, device=X.device)
        _lib.check(_lib.load().arcq_absmax(X.data_ptr(), X.numel(), out.data_ptr(), _stream(X)), "arcq_absmax")
    return out


def _c_reorder_quantize_x_scaled(X, reorder_index, KE, flavour: int = -1, amax=None, mode: int = 0):
    if amax is None:
        amax = _c_absmax(X)
    else:
        _need(amax, "amax", torch.float32)
        if amax.numel() != 1 or amax.device != X.device:
            raise RuntimeError("amax must be one float32 on %s" % X.device)
    return _quant("arcq_quantize_x_scaled", X, reorder_index, KE, flavour, amax, mode)


def _c_rmsnorm_quantize_x(X, W, eps, reorder_index, KE):
    """Rmsnorm and quantize activation, return K + KE channels (bindings.cpp:216-254)."""
    L = _lib.load()
    _need(X, "X", torch.bfloat16, 2)
    _need(W, "W", torch.bfloat16, 1)
    _need(reorder_index, "reorder_index", torch.int16, 1)
    M, H = X.shape
    if W.numel() != H or reorder_index.numel() != H:
        raise RuntimeError("W and reorder_index must have %d entries" % H)
    if W.device != X.device or reorder_index.device != X.device:
        raise RuntimeError("W and reorder_index must be on %s" % X.device)
    KE = int(KE)
    K = H + KE
    with torch.cuda.device(X.device):
        q = torch.empty((M, K // 2), dtype=torch.uint8, device=X.device)
        sf = torch.empty((_sf_bytes(M, K),), dtype=torch.uint8, device=X.device)
        rc = L.arcq_rmsnorm_quantize_x(X.data_ptr(), W.data_ptr(), float(eps), M, H, reorder_index.data_ptr(),
                                       KE, q.data_ptr(), sf.data_ptr(), _stream(X))
    _lib.check(rc, "arcq_rmsnorm_quantize_x")
    return q, sf


def _c_add_rmsnorm_quantize_x(X, residual, W, eps, reorder_index, KE):
    """residual + hidden_states, then the norm + quantise of the next block, in one launch (modeling_arc.py:294-301):
    returns (QX, SFX, X + residual)."""
    L = _lib.load()
    _need(X, "X", torch.bfloat16, 2)
    _need(residual, "residual", torch.bfloat16, 2)
    _need(W, "W", torch.bfloat16, 1)
    _need(reorder_index, "reorder_index", torch.int16, 1)
    M, H = X.shape
    if tuple(residual.shape) != (M, H) or W.numel() != H or reorder_index.numel() != H:
        raise RuntimeError("residual must have the shape of X; W and reorder_index must have %d entries" % H)
    if W.device != X.device or reorder_index.device != X.device or residual.device != X.device:
        raise RuntimeError("all inputs must be on %s" % X.device)
    KE = int(KE)
    K = H + KE
    with torch.cuda.device(X.device):
        q = torch.empty((M, K // 2), dtype=torch.uint8, device=X.device)
        sf = torch.empty((_sf_bytes(M, K),), dtype=torch.uint8, device=X.device)
        s = torch.empty((M, H), dtype=torch.bfloat16, device=X.device)
        rc = L.arcq_add_rmsnorm_quantize_x(X.data_ptr(), residual.data_ptr(), W.data_ptr(), float(eps), M, H, reorder_index.data_ptr(),
                                           KE, s.data_ptr(), q.data_ptr(), sf.data_ptr(), _stream(X))
    _lib.check(rc, "arcq_add_rmsnorm_quantize_x")
    return q, sf, s


def _c_matmul(A, B, SFA, SFB, scale, col_scale=None, epilogue: int = 0, out=None, bias=None):
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
    if B.device != A.device or SFA.device != A.device or SFB.device != A.device:
        raise RuntimeError("B, SFA and SFB must be on %s" % A.device)
    K = 2 * K2
    need_a = ((M + 127) // 128) * 128 * K // 16
    need_b = ((N + 127) // 128) * 128 * K // 16
    if SFA.numel() < need_a or SFB.numel() < need_b:
        raise RuntimeError("scale-factor buffers too small for M=%d N=%d K=%d" % (M, N, K))
    alpha_dev = 0
    alpha = 0.0
    if (isinstance(scale, torch.Tensor) and scale.is_cuda and scale.dtype == torch.float32 and scale.numel() == 1
            and scale.device == A.device):
        alpha_dev = scale.data_ptr()  # read by the kernel: no host sync
    else:
        alpha = float(scale)  # Python float, or any other tensor through .item() like the reference's `const float scale`
    cs = 0
    if col_scale is not None:
        _need(col_scale, "col_scale", torch.float32, 1)
        if col_scale.numel() != N or col_scale.device != A.device:
            raise RuntimeError("col_scale must hold %d float32 on %s" % (N, A.device))
        cs = col_scale.data_ptr()
    bp = 0
    if bias is not None:  # y + self.bias of the reference's linear (qLinearLayer.py:74-75), in the epilogue
        _need(bias, "bias", torch.bfloat16, 1)
        if bias.numel() != N or bias.device != A.device or epilogue != EPILOGUE_NONE:
            raise RuntimeError("bias must hold %d bfloat16 on %s and needs epilogue == EPILOGUE_NONE" % (N, A.device))
        bp = bias.data_ptr()
    out_cols = N // 2 if epilogue == EPILOGUE_SWIGLU else N
    with torch.cuda.device(A.device):
        if out is None:
            out = torch.empty((M, out_cols), dtype=torch.bfloat16, device=A.device)
        else:
            if not (isinstance(out, torch.Tensor) and out.is_cuda and out.device == A.device and out.dtype == torch.bfloat16
                    and out.dim() == 2):
                raise RuntimeError("out must be a 2-D bf16 tensor on %s" % A.device)
            if tuple(out.shape) != (M, out_cols) or out.stride(1) != 1 or out.stride(0) < out_cols:
                raise RuntimeError("out must be [%d, %d] with unit column stride, got %s strides %s"
                                   % (M, out_cols, tuple(out.shape), out.stride()))
        rc = L.arcq_gemm_nvfp4_ex2(A.data_ptr(), B.data_ptr(), SFA.data_ptr(), SFB.data_ptr(), M, N, K, alpha,
                                   alpha_dev, cs, bp, int(epilogue), out.data_ptr(), out.stride(0), _stream(A))
    _lib.check(rc, "arcq_gemm_nvfp4")
    return out


def _not_on_this_path(name):
    def f(*a, **k):
        raise NotImplementedError(
            "agemm.%s is the reference's int4 KV-cache path (bindings.cpp:258-543), outside the NVFP4+ARC hot path" % name)
    f.__name__ = name
    return f


_KV = ("batch_decode_i4", "init_kv_i4", "append_kv_i4", "batch_decode_f16", "init_kv_f16", "append_kv_f16")

# ---- bind the public names: compiled module first ------------------------------------------------------------------

_ext = _lib.load_ext()
if _ext is not None:
    BACKEND = "pybind11"
    matmul = _ext.matmul
    reorder_quantize_x = _ext.reorder_quantize_x
    reorder_quantize_w = _ext.reorder_quantize_w
    rmsnorm_quantize_x = _ext.rmsnorm_quantize_x
    add_rmsnorm_quantize_x = _ext.add_rmsnorm_quantize_x
    reorder_quantize_x_scaled = _ext.reorder_quantize_x_scaled
    absmax = _ext.absmax
    sf_bytes = _ext.sf_bytes
    flavour_for_kq = _ext.flavour_for_kq
    launch_count = _ext.launch_count
    set_option = _ext.set_option
    use_plans = _ext.use_plans
    plan_build = _ext.plan_build
    plan_max_conflict = _ext.plan_max_conflict
    reorder_quantize_x_planned = _ext.reorder_quantize_x_planned
    for _n in _KV:
        globals()[_n] = getattr(_ext, _n)
else:
    BACKEND = "ctypes"
    matmul = _c_matmul
    reorder_quantize_x = _c_reorder_quantize_x
    reorder_quantize_w = _c_reorder_quantize_w
    rmsnorm_quantize_x = _c_rmsnorm_quantize_x
    add_rmsnorm_quantize_x = _c_add_rmsnorm_quantize_x
    reorder_quantize_x_scaled = _c_reorder_quantize_x_scaled
    absmax = _c_absmax
    sf_bytes = _c_sf_bytes
    flavour_for_kq = _c_flavour_for_kq
    launch_count = _c_launch_count
    set_option = _c_set_option

    def use_plans(on):  # the ctypes binding never builds placement plans
        return None

    for _n in _KV:
        globals()[_n] = _not_on_this_path(_n)


# ---- weight packing for the fused launches (pure tensor bookkeeping on already-quantised weights) ------------------

def _sf_blocks(sf: torch.Tensor, rows: int, K: int) -> torch.Tensor:
    """The valid part of a swizzled scale buffer as [row blocks of 128, K/64 atoms x 512 bytes]."""
    nblk = (rows + 127) // 128
    return sf[: nblk * 128 * K // 16].view(nblk, (K // 64) * 512)


def fuse_weights(parts):
    """Concatenate quantised weights that share one quantised input (q/k/v: qLlamaLayer.py:301-303) along N.

    parts: list of (QW u8 [N_i, K/2], SFW u8 swizzled, scale_w float | 0-dim/1-elem tensor); every N_i must be a multiple
    of 128 so that the 128-row scale-factor blocks concatenate without touching a byte.  Returns (QW [sum N_i, K/2],
    SFW, col_scale fp32 [sum N_i]) for ``matmul(..., scale_x, col_scale=col_scale)``: column n is scaled by
    RN(scale_x * scale_w[n]) -- the same fp32 product the reference forms per layer (qLinearLayer.py:69).
    """
    K2 = parts[0][0].shape[1]
    dev = parts[0][0].device
    qs, sfs, cs = [], [], []
    for qw, sfw, sw in parts:
        n = qw.shape[0]
        if n % 128 or qw.shape[1] != K2:
            raise RuntimeError("fuse_weights: every part needs N %% 128 == 0 and the same K")
        qs.append(qw)
        sfs.append(_sf_blocks(sfw, n, 2 * K2))
        sw_t = sw.reshape(1).to(device=dev, dtype=torch.float32) if isinstance(sw, torch.Tensor) else \
            torch.tensor([float(sw)], dtype=torch.float32, device=dev)
        cs.append(sw_t.expand(n))
    N = sum(q.shape[0] for q in qs)
    sf = torch.zeros((_sf_bytes(N, 2 * K2),), dtype=torch.uint8, device=dev)
    body = torch.cat(sfs, 0).reshape(-1)
    sf[: body.numel()] = body
    return torch.cat(qs, 0).contiguous(), sf, torch.cat(cs).contiguous()


def interleave_gate_up(gate, up):
    """Fused gate/up weight for ``matmul(..., epilogue=EPILOGUE_SWIGLU)`` (modeling_arc.py:109-111, qLlamaLayer.py:432).

    gate, up: (QW u8 [I, K/2], SFW, scale_w) with I a multiple of 128.  Rows are interleaved in 128-row blocks --
    [gate block 0, up block 0, gate block 1, ...] -- so that one 256-wide GEMM tile holds the gate and the up columns of
    the same 128 channels.  Returns (QW [2I, K/2], SFW, col_scale fp32 [2I]).
    """
    (qg, sg, wg), (qu, su, wu) = gate, up
    I, K2 = qg.shape
    if qu.shape != qg.shape or I % 128:
        raise RuntimeError("interleave_gate_up: gate and up must both be [I, K/2] with I %% 128 == 0")
    dev = qg.device
    nb = I // 128
    q = torch.stack((qg.view(nb, 128, K2), qu.view(nb, 128, K2)), 1).reshape(2 * I, K2).contiguous()
    K = 2 * K2
    body = torch.stack((_sf_blocks(sg, I, K), _sf_blocks(su, I, K)), 1).reshape(-1)
    sf = torch.zeros((_sf_bytes(2 * I, K),), dtype=torch.uint8, device=dev)
    sf[: body.numel()] = body

    def vec(s):
        t = s.reshape(1).to(device=dev, dtype=torch.float32) if isinstance(s, torch.Tensor) else \
            torch.tensor([float(s)], dtype=torch.float32, device=dev)
        return t.expand(nb, 128)

    cs = torch.stack((vec(wg), vec(wu)), 1).reshape(-1).contiguous()
    return q, sf, cs
