"""CPU oracle for the NVFP4 + ARC hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker (or as
the timed CPU baseline).  The product (``arcquant_b200``/``agemm``) never does.

``oracle.lib()`` returns numpy-level wrappers around ``arcq_oracle.c`` (the
kernel-faithful restatement, see the header of that file for what pins it).
``oracle.fakequant`` is a torch restatement of the reference's PyTorch fake-quant
path (``model/quantize.py``), the reference's only CPU-runnable implementation
and therefore the timed CPU baseline.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libarcq_oracle.so")
_SRC = os.path.join(_HERE, "arcq_oracle.c")


def build(force: bool = False) -> str:
    """Compile arcq_oracle.c with gcc (no FMA contraction)."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        tmp = _SO + ".tmp%d" % os.getpid()
        subprocess.check_call(
            ["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", tmp, _SRC, "-lm"]
        )
        os.replace(tmp, _SO)
    return _SO


def _u8p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))


def _u16p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint16))


def _i16p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int16))


class Oracle:
    """numpy front-end.  bf16 tensors travel as uint16 arrays (raw bits)."""

    def __init__(self):
        L = ctypes.CDLL(build())
        self.L = L
        L.arcq_o_sf_offset.restype = ctypes.c_int64
        L.arcq_o_sf_offset.argtypes = [ctypes.c_int64] * 3
        L.arcq_o_sf_bytes.restype = ctypes.c_int64
        L.arcq_o_sf_bytes.argtypes = [ctypes.c_int64] * 2
        L.arcq_o_flavour_for_kq.restype = ctypes.c_int
        L.arcq_o_flavour_for_kq.argtypes = [ctypes.c_int]
        L.arcq_o_bf16_to_f32.restype = ctypes.c_float
        L.arcq_o_bf16_to_f32.argtypes = [ctypes.c_uint16]
        L.arcq_o_f32_to_bf16.restype = ctypes.c_uint16
        L.arcq_o_f32_to_bf16.argtypes = [ctypes.c_float]
        L.arcq_o_f32_to_ue4m3.restype = ctypes.c_uint8
        L.arcq_o_f32_to_ue4m3.argtypes = [ctypes.c_float]
        L.arcq_o_ue4m3_to_f32.restype = ctypes.c_float
        L.arcq_o_ue4m3_to_f32.argtypes = [ctypes.c_uint8]
        L.arcq_o_f32_to_e2m1.restype = ctypes.c_uint8
        L.arcq_o_f32_to_e2m1.argtypes = [ctypes.c_float]
        L.arcq_o_e2m1_to_f32.restype = ctypes.c_float
        L.arcq_o_e2m1_to_f32.argtypes = [ctypes.c_uint8]
        L.arcq_o_rms_sumsq.restype = ctypes.c_float
        L.arcq_o_rms_sumsq.argtypes = [ctypes.POINTER(ctypes.c_uint16), ctypes.c_int]
        P8, P16, PI = (ctypes.POINTER(ctypes.c_uint8), ctypes.POINTER(ctypes.c_uint16),
                       ctypes.POINTER(ctypes.c_int16))
        for f in (L.arcq_o_quantize_x, L.arcq_o_quantize_w):
            f.restype = ctypes.c_int
            f.argtypes = [P16, ctypes.c_int64, ctypes.c_int, PI, ctypes.c_int, ctypes.c_int, P8, P8]
        L.arcq_o_rmsnorm_quantize_x.restype = ctypes.c_int
        L.arcq_o_rmsnorm_quantize_x.argtypes = [
            P16, P16, ctypes.c_float, ctypes.c_int64, ctypes.c_int, PI, ctypes.c_int,
            ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float), P8, P8]
        L.arcq_o_dequant.restype = None
        L.arcq_o_dequant.argtypes = [P8, P8, ctypes.c_int64, ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
        L.arcq_o_gemm.restype = ctypes.c_int
        L.arcq_o_gemm.argtypes = [P8, P8, P8, P8, ctypes.c_int64, ctypes.c_int64, ctypes.c_int,
                                  ctypes.c_float, ctypes.POINTER(ctypes.c_double)]

    # -- layout -----------------------------------------------------------
    def sf_offset(self, row, g, K):
        return int(self.L.arcq_o_sf_offset(row, g, K))

    def sf_bytes(self, rows, K):
        return int(self.L.arcq_o_sf_bytes(rows, K))

    def flavour_for_kq(self, KQ):
        return int(self.L.arcq_o_flavour_for_kq(KQ))

    def sf_valid_offsets(self, rows, K):
        """All byte offsets the kernels write: (rows x K/16) int64 array."""
        r = np.arange(rows, dtype=np.int64)[:, None]
        g = np.arange(K // 16, dtype=np.int64)[None, :]
        return (r // 128) * (K // 64) * 512 + (g // 4) * 512 + (r % 32) * 16 + ((r // 32) % 4) * 4 + (g % 4)

    # -- kernels ----------------------------------------------------------
    def _quant(self, fn, x, idx, KE, flavour):
        x = np.ascontiguousarray(x, dtype=np.uint16)
        idx = np.ascontiguousarray(idx, dtype=np.int16)
        M, KQ = x.shape
        if flavour is None:
            flavour = max(self.flavour_for_kq(KQ), 0)
        K = KQ + KE
        q = np.zeros((M, K // 2), dtype=np.uint8)
        sf = np.zeros(self.sf_bytes(M, K), dtype=np.uint8)
        rc = fn(_u16p(x), M, KQ, _i16p(idx), KE, flavour, _u8p(q), _u8p(sf))
        if rc != 0:
            raise ValueError("oracle: unsupported shape KQ=%d KE=%d flavour=%d" % (KQ, KE, flavour))
        return q, sf

    def quantize_x(self, x_bits, idx, KE, flavour=None):
        return self._quant(self.L.arcq_o_quantize_x, x_bits, idx, KE, flavour)

    def quantize_w(self, w_bits, idx, KE, flavour=None):
        return self._quant(self.L.arcq_o_quantize_w, w_bits, idx, KE, flavour)

    def rmsnorm_quantize_x(self, x_bits, w_bits, eps, idx, KE, rstd=None):
        x = np.ascontiguousarray(x_bits, dtype=np.uint16)
        w = np.ascontiguousarray(w_bits, dtype=np.uint16)
        idx = np.ascontiguousarray(idx, dtype=np.int16)
        M, H = x.shape
        K = H + KE
        q = np.zeros((M, K // 2), dtype=np.uint8)
        sf = np.zeros(self.sf_bytes(M, K), dtype=np.uint8)
        rout = np.zeros(M, dtype=np.float32)
        rin = None
        if rstd is not None:
            rstd = np.ascontiguousarray(rstd, dtype=np.float32)
            rin = rstd.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
        rc = self.L.arcq_o_rmsnorm_quantize_x(
            _u16p(x), _u16p(w), ctypes.c_float(eps), M, H, _i16p(idx), KE, rin,
            rout.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), _u8p(q), _u8p(sf))
        if rc != 0:
            raise ValueError("oracle: unsupported rmsnorm shape H=%d KE=%d" % (H, KE))
        return q, sf, rout

    def rms_sumsq(self, row_bits):
        row = np.ascontiguousarray(row_bits, dtype=np.uint16)
        return float(self.L.arcq_o_rms_sumsq(_u16p(row), row.shape[0]))

    def dequant(self, q, sf, K):
        q = np.ascontiguousarray(q, dtype=np.uint8)
        sf = np.ascontiguousarray(sf, dtype=np.uint8)
        rows = q.shape[0]
        out = np.zeros((rows, K), dtype=np.float32)
        self.L.arcq_o_dequant(_u8p(q), _u8p(sf), rows, K, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
        return out

    def gemm(self, a, b, sfa, sfb, alpha):
        """float64 dequantise-then-matmul reference (small shapes only)."""
        a = np.ascontiguousarray(a, dtype=np.uint8)
        b = np.ascontiguousarray(b, dtype=np.uint8)
        M, K = a.shape[0], a.shape[1] * 2
        N = b.shape[0]
        out = np.zeros((M, N), dtype=np.float64)
        rc = self.L.arcq_o_gemm(_u8p(a), _u8p(b), _u8p(np.ascontiguousarray(sfa)),
                                _u8p(np.ascontiguousarray(sfb)), M, N, K, ctypes.c_float(alpha),
                                out.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
        if rc != 0:
            raise ValueError("oracle: K must be a multiple of 64")
        return out

    def gemm_fast(self, a, b, sfa, sfb, alpha):
        """Same reference through a float64 numpy matmul of the dequantised operands."""
        K = a.shape[1] * 2
        da = self.dequant(a, sfa, K).astype(np.float64)
        db = self.dequant(b, sfb, K).astype(np.float64)
        return float(np.float32(alpha)) * (da @ db.T)


_LIB = None


def lib() -> Oracle:
    global _LIB
    if _LIB is None:
        _LIB = Oracle()
    return _LIB


# ---- numpy bf16 helpers (bit-level, shared by tests and bench) -----------------

def f32_to_bf16_bits(a: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even float32 -> bf16 bit pattern (uint16)."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    lsb = (u >> 16) & 1
    return ((u + 0x7FFF + lsb) >> 16).astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)
