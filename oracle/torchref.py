"""float64 dequantise-then-matmul oracle on SAMPLED rows, in plain torch -- TEST INFRASTRUCTURE ONLY.

Same semantics as ``arcq_o_dequant`` / ``arcq_o_gemm`` of arcq_oracle.c (reference: kernels/src/nvfp4.cu:98-121,
SURVEY.md Appendix A.7), written with torch ops so that it runs on whatever device holds the operands: the bench
shapes (M = 8192, N up to 28672, K up to 14848) are far too big for the C oracle, but a float64 cuBLAS matmul of a few
hundred sampled rows against the whole dequantised weight is cheap and shares no code with the kernels under test.
``tests/test_oracle.py`` pins it against the C oracle on the CPU.  Only tests/ and bench.py's post-run ``--check`` leg
import it.
"""
from __future__ import annotations

import torch

_E2M1 = [0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0]


def _luts(device):
    e2m1 = torch.tensor(_E2M1 + [-v for v in _E2M1], dtype=torch.float64, device=device)  # sign | e e | m
    b = torch.arange(256, dtype=torch.int64, device=device)
    e, m = (b >> 3) & 15, (b & 7).double()
    ue4m3 = torch.where(e == 0, m / 8.0 * 2.0 ** -6, (1.0 + m / 8.0) * torch.pow(2.0, (e - 7).double()))  # bias 7, subnormals
    return e2m1, ue4m3


def sf_offsets(rows: torch.Tensor, K: int) -> torch.Tensor:
    """Byte offset of the scale of (row, 16-group): the 128x4 swizzle (reorder.cu:140-143; arcq_sf_offset)."""
    r = rows.to(torch.int64)[:, None]
    g = torch.arange(K // 16, dtype=torch.int64, device=rows.device)[None, :]
    return (r // 128) * (K // 64) * 512 + (g // 4) * 512 + (r % 32) * 16 + ((r // 32) % 4) * 4 + (g % 4)


def dequant_rows(q: torch.Tensor, sf: torch.Tensor, rows: torch.Tensor, K: int) -> torch.Tensor:
    """float64 [len(rows), K] = e2m1 value * E4M3 scale of the selected rows of (q u8 [R, K/2], sf swizzled u8)."""
    e2m1, ue4m3 = _luts(q.device)
    rows = rows.to(q.device)
    qs = q[rows].to(torch.int64)
    codes = torch.stack((qs & 15, qs >> 4), -1).reshape(rows.numel(), K)  # low nibble = even element
    vals = e2m1[codes]
    scales = ue4m3[sf[sf_offsets(rows, K)].to(torch.int64)]
    return (vals.view(rows.numel(), K // 16, 16) * scales[:, :, None]).reshape(rows.numel(), K)


def gemm_rows(qa, sfa, qb, sfb, rows, alpha, col_scale=None, chunk=4096):
    """(ref, mag) float64 [len(rows), N]: alpha * A[rows] . B^T and alpha * |A[rows]| . |B|^T (the tolerance needs both).
    alpha is rounded to fp32 first, then (if given) multiplied per column in fp32 like the kernels do."""
    K = qa.shape[1] * 2
    N = qb.shape[0]
    da = dequant_rows(qa, sfa, rows, K)
    ref = torch.empty((rows.numel(), N), dtype=torch.float64, device=qa.device)
    mag = torch.empty_like(ref)
    for n0 in range(0, N, chunk):
        nr = torch.arange(n0, min(N, n0 + chunk), device=qa.device)
        db = dequant_rows(qb, sfb, nr, K)
        ref[:, n0:n0 + nr.numel()] = da @ db.T
        mag[:, n0:n0 + nr.numel()] = da.abs() @ db.abs().T
    a32 = torch.tensor(float(alpha), dtype=torch.float32, device=qa.device)
    if col_scale is not None:
        a = (a32 * col_scale.to(torch.float32)).double()[None, :]
    else:
        a = a32.double()
    return ref * a, mag * a.abs()


def sample_rows(M: int, per_block: int = 2, seed: int = 0) -> torch.Tensor:
    """`per_block` rows of every 128-row block (so every M tile, every CTA of a pair and every wave is hit), plus the
    first and last row."""
    g = torch.Generator().manual_seed(seed)
    out = [0, M - 1]
    for b in range((M + 127) // 128):
        lo, hi = b * 128, min(M, b * 128 + 128)
        out += (lo + torch.randint(0, hi - lo, (per_block,), generator=g)).tolist()
    return torch.tensor(sorted(set(out)), dtype=torch.int64)


def check_gemm(y, qa, sfa, qb, sfb, alpha, rows=None, col_scale=None):
    """Worst |y - ref| / tol over the sampled rows and the relative Frobenius error; tol = 2^-8 |ref| + 2^-18 mag
    (tests/test_gpu_gemm.py: one bf16 rounding + fp32 accumulation-order noise)."""
    M = y.shape[0]
    if rows is None:
        rows = sample_rows(M)
    rows = rows.to(y.device)
    ref, mag = gemm_rows(qa, sfa, qb, sfb, rows, alpha, col_scale)
    got = y[rows].double()
    err = (got - ref).abs()
    tol = 2.0 ** -8 * ref.abs() + 2.0 ** -18 * mag + 1e-30
    worst = float((err / tol).max())
    rel = float((got - ref).norm() / ref.norm().clamp_min(1e-30))
    return worst, rel
