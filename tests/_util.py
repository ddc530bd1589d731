"""Synthetic inputs shared by the parity tests (SURVEY.md §8d "Synthetic inputs")."""
from __future__ import annotations

import numpy as np
import torch


def make_acts(M, KQ, KE, seed=0, dist="A"):
    """bf16 activations [M, KQ] + ascending-absmax reorder index (int16), pre-scaled like the callers do
    (x / (max|x| / 2688) in bf16; reference model/qLlamaLayer.py:73-77)."""
    g = torch.Generator().manual_seed(seed)
    if dist == "A":  # N(0,1) body, KE heavy-tailed outlier channels at random positions
        x = torch.randn(M, KQ, generator=g)
        perm = torch.randperm(KQ, generator=g)
        if KE:
            x[:, perm[:KE]] *= 20.0
    elif dist == "B":  # the reference's kernels/main.py:13-19 distribution
        x = torch.rand(M, KQ, generator=g) * 3
        x[:, -384:] += 3
        x[:, -128:] = x[:, -128:] * 8 + 8
        x[:, -16:] = x[:, -16:] * 32 + 32
        sign = torch.randint(0, 2, (M, KQ), generator=g) * 2 - 1
        x = x * sign
        perm = torch.randperm(KQ, generator=g)
        x = x[:, perm]
    elif dist == "small":  # tiny magnitudes: exercises E4M3 subnormal scales and the 2^-9 clamp
        x = torch.randn(M, KQ, generator=g) * 1e-3
        x[:, ::7] = 0.0
    else:
        raise ValueError(dist)
    x = x.to(torch.bfloat16)
    if dist != "small":
        scale = x.abs().max().float() / (448.0 * 6.0)
        x = (x / scale).to(torch.bfloat16)
    amax = x.float().abs().amax(dim=0)
    idx = torch.argsort(amax, stable=True).to(torch.int16)
    return x.contiguous(), idx.contiguous()


def make_weights(N, KQ, seed=1):
    g = torch.Generator().manual_seed(seed)
    w = (torch.randn(N, KQ, generator=g) * 0.02).to(torch.bfloat16)
    scale = w.max().float() / (448.0 * 6.0)  # signed max: reference model/qLinearLayer.py:26
    return (w / scale).to(torch.bfloat16).contiguous(), float(scale)


def bits(t: torch.Tensor) -> np.ndarray:
    """bf16 tensor -> uint16 bit pattern array."""
    return t.contiguous().view(torch.int16).cpu().numpy().view(np.uint16)


def sf_valid(oracle, sf: np.ndarray, rows: int, K: int) -> np.ndarray:
    """The scale bytes the kernels define: [rows, K/16] gathered from the swizzled buffer."""
    return np.asarray(sf)[oracle.sf_valid_offsets(rows, K)]
