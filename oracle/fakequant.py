"""Torch restatement of the reference's PyTorch fake-quant path -- TEST INFRASTRUCTURE / CPU BASELINE ONLY.

Follows /root/reference/model/quantize.py: quantize_e2m1 (:14-22), quantize_ue4m3 (:40-49),
quantize_nvfp4_tensor (:65-92), fake_reorder_quantize_w (:219-241), fake_reorder_quantize_x (:243-268).
It is the reference's only CPU-runnable implementation of the hot path and therefore the timed CPU
baseline (bench.py `cpu_baseline` / `--impl reference`); it is pinned against the real module by
tests/golden/make_golden.py (run where /root/reference exists) + tests/test_oracle.py.

Known, deliberate differences of this path from the CUDA kernels: e2m1 ties go to
the first minimum of the value table, no -0, no E4M3 subnormals (clamp at 2e-3), true division, and the
primary groups are formed in ORIGINAL channel order (only the residual uses reorder_index[-KE:]).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

_E2M1 = [-6.0, -4.0, -3.0, -2.0, -1.5, -1.0, -0.5, 0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0]


def quantize_e2m1(t: torch.Tensor) -> torch.Tensor:  # quantize.py:14-22
    vals = torch.tensor(_E2M1, device=t.device, dtype=t.dtype)
    diff = torch.abs(t.unsqueeze(-1) - vals)
    return vals[torch.argmin(diff, dim=-1)]


def quantize_ue4m3(t: torch.Tensor) -> torch.Tensor:  # quantize.py:40-49
    t = torch.clamp(t, min=2e-3, max=448.0)
    exponent = torch.floor(torch.log2(t + 1e-9))
    mant = t / (2 ** exponent) - 1.0
    qmant = torch.round(mant * 8) / 8
    return (1 + qmant) * (2 ** exponent)


def quantize_nvfp4_tensor(t: torch.Tensor, group_size: int = 16) -> torch.Tensor:  # quantize.py:65-92
    shape = t.shape
    pad = (group_size - t.shape[-1] % group_size) % group_size
    if pad:
        t = F.pad(t, (0, pad))
    g = t.view(-1, group_size) if t.is_contiguous() else t.reshape(-1, group_size)
    amax = torch.max(torch.abs(g), dim=1, keepdim=True)[0]
    scale = amax / 6.0
    scale[scale == 0] = 1e-9
    dscale = quantize_ue4m3(scale)
    deq = (quantize_e2m1(g / dscale) * dscale).view(t.shape)
    if pad:
        deq = deq[..., :-pad]
    return deq.view(shape) if deq.is_contiguous() else deq.reshape(shape)


def fake_reorder_quantize_w(w, reorder_index, select_num):  # quantize.py:219-241 (NVFP4 branch)
    scale = torch.max(w).float() / (448.0 * 6.0)
    w = w / scale
    scale_w = w.abs().max(dim=1, keepdim=True)[0]
    if select_num == 0:
        return quantize_nvfp4_tensor(w), scale_w, scale
    topk = reorder_index[-select_num:]
    return torch.cat([quantize_nvfp4_tensor(w), quantize_nvfp4_tensor(w[:, topk])], dim=1), scale_w, scale


def fake_reorder_quantize_x(x, reorder_index, select_num):  # quantize.py:243-268 (NVFP4 branch)
    scale = torch.max(x).float() / (448.0 * 6.0)
    x = x / scale
    scale_x = x.abs().max(dim=1, keepdim=True)[0]
    if select_num == 0:
        return quantize_nvfp4_tensor(x), scale_x, scale
    topk = reorder_index[-select_num:]
    q_x = quantize_nvfp4_tensor(x)
    err = x - q_x
    q_err = quantize_nvfp4_tensor(err[:, topk])
    return torch.cat([q_x, q_err], dim=1) * scale, scale_x, scale


def fake_linear(x, w, reorder_index, select_num):
    """The reference's fake-quant linear: quantise both sides, then one dense matmul over K + KE."""
    qx, _, _ = fake_reorder_quantize_x(x, reorder_index, select_num)
    qw, _, sw = fake_reorder_quantize_w(w, reorder_index, select_num)
    return F.linear(qx, qw) * sw
