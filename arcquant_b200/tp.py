"""Tensor-parallel sharding of NVFP4+ARC linears -- host logic + thin layers over ``agemm``.

The reference has no distributed code; its single-GPU layer is ``QLinearLayer`` (model/qLinearLayer.py:29-77) fed by
``NVFP4_reorder_quantize_x`` (model/qLlamaLayer.py:73-77).  Linear layers shard naturally:

* column-parallel (q/k/v, gate/up): split N.  Every rank quantises the replicated activation with the FULL
  ``reorder_index`` / ``select_num`` and multiplies by its ``[N/R, K]`` weight slice.  No collective.
* row-parallel (o, down): split K by ORIGINAL channel ownership (contiguous slices: whole heads for o, the rank's
  own gate/up columns for down).  Rank r keeps the global order restricted to its channels (``shard_reorder_index``)
  and ``KE_r`` = the number of the global top-``select_num`` channels it owns, rounded up to a multiple of 64 (so the
  ARC residual columns stay with their owning shard).  Partial ``[M, N]`` results are summed with one all-reduce
  (NCCL over NVLink on GPUs; gloo in the CPU tests).

Nothing here computes on the CPU: the layers call the CUDA kernels through ``agemm``; the pure index arithmetic
(`shard_*`, `split_rows`) is what the world_size-2 gloo tests exercise.
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist

ARC_ALIGN = 64  # select_num granularity of the reference (utilize.py:475); also one scale-factor atom column


def split_rows(n: int, world: int, rank: int, align: int = 1) -> Tuple[int, int]:
    """[start, stop) of rank's contiguous slice of n rows/channels; slice boundaries are multiples of `align`."""
    if n % (world * align):
        raise ValueError("cannot split %d into %d slices aligned to %d" % (n, world, align))
    per = n // world
    return rank * per, (rank + 1) * per


def shard_reorder_index(reorder_index: torch.Tensor, select_num: int, world: int, rank: int,
                        align: int = ARC_ALIGN) -> Tuple[torch.Tensor, int]:
    """Row-parallel shard of a reorder permutation.

    Returns (local int16 permutation of rank's KQ/world channels, KE_r).  The local order is the global order
    restricted to the rank's contiguous channel slice [lo, hi) (stable), re-based to local offsets; KE_r counts the
    global top-`select_num` channels (the LAST select_num entries of reorder_index, utilize.py:42) the rank owns,
    rounded up to `align` and capped at the shard width.
    """
    idx = reorder_index.to(torch.int64).cpu()
    KQ = idx.numel()
    lo, hi = split_rows(KQ, world, rank, 64)
    mine = (idx >= lo) & (idx < hi)
    local = (idx[mine] - lo).to(torch.int16)
    owned = int(mine[KQ - select_num:].sum()) if select_num > 0 else 0
    ke = min(-(-owned // align) * align, hi - lo)
    return local.contiguous(), ke


def shard_select_nums(reorder_index: torch.Tensor, select_num: int, world: int) -> List[int]:
    return [shard_reorder_index(reorder_index, select_num, world, r)[1] for r in range(world)]


def nvfp4_scale(t: torch.Tensor, signed: bool) -> torch.Tensor:
    """Per-tensor pre-scale of the reference callers: max|x| / 2688 for activations (qLlamaLayer.py:73-77), signed
    max(w) / 2688 for weights (qLinearLayer.py:25-28).  Stays on the device (no host sync)."""
    m = t.max() if signed else t.abs().max()
    return m.float() / (448.0 * 6.0)


class QLinearShard(torch.nn.Module):
    """One rank's slice of a QLinearLayer (model/qLinearLayer.py:29-77): quantised weight + per-tensor scale."""

    def __init__(self, weight: torch.Tensor, reorder_index: torch.Tensor, select_num: int, bias=None, flavour: int = -1):
        """flavour: ARC interleave / residual formula of BOTH sides of this linear (-1 = the reference's KQ table).
        Layers fed by ``rmsnorm_quantize_x`` must pass FLAVOUR_16: that entry point is always 16-wide (bindings.cpp:243)
        while the KQ table picks the 32-wide weight layout for KQ in {3584, 18944, 27648, 28672} (:199) -- the
        reference's own mismatch, which mis-pairs the ARC columns of X and W whenever KE > 0."""
        super().__init__()
        from . import agemm
        self.agemm = agemm
        self.out_features, self.in_features = weight.shape
        self.select_num = int(select_num)
        self.flavour = int(flavour)
        self.register_buffer("reorder_index", reorder_index.to(torch.int16).to(weight.device).contiguous())
        scale = nvfp4_scale(weight, signed=True)
        w = (weight / scale).to(torch.bfloat16).contiguous()
        self.W, self.scale_w = agemm.reorder_quantize_w(w, self.reorder_index, self.select_num, self.flavour)
        self.register_buffer("scale", scale.reshape(1))
        self.bias = bias

    @torch.no_grad()
    def quantize_input(self, x: torch.Tensor, fused: bool = True):
        """NVFP4_reorder_quantize_x (qLlamaLayer.py:73-77) without the host sync: the scale stays a device scalar.
        fused: max|x| pre-pass + in-kernel `bf16(x / scale)` (two launches, X read twice) instead of torch's abs / max /
        divide passes in front of the kernel; same bytes (tests/test_gpu_prescale.py)."""
        if fused and self.in_features % 64 == 0 and self.select_num % 64 == 0:  # shapes of the TMA-staged kernel
            return self.agemm.reorder_quantize_x_scaled(x, self.reorder_index, self.select_num, self.flavour)
        scale = nvfp4_scale(x, signed=False).reshape(1)
        qx, scale_x = self.agemm.reorder_quantize_x((x / scale).to(torch.bfloat16), self.reorder_index, self.select_num,
                                                    self.flavour)
        return qx, scale_x, scale

    @torch.no_grad()
    def forward_quantized(self, qx, scale_x, scale, out=None):
        b = self.bias
        if b is not None and b.dtype == torch.bfloat16 and b.is_contiguous() and b.data_ptr() % 16 == 0:
            # `y + self.bias` (qLinearLayer.py:74-75) in the GEMM epilogue: same bits, no extra pass over [M, N]
            return self.agemm.matmul(qx, self.W, scale_x, self.scale_w, scale * self.scale, out=out, bias=b)
        y = self.agemm.matmul(qx, self.W, scale_x, self.scale_w, scale * self.scale, out=out)
        if b is not None:
            y = y + b
        return y

    @torch.no_grad()
    def forward(self, x: torch.Tensor):
        return self.forward_quantized(*self.quantize_input(x))


class ColumnParallelQLinear(QLinearShard):
    """N-split: this rank holds rows [lo, hi) of the weight; input replicated; output is the rank's N/world columns."""

    def __init__(self, weight, reorder_index, select_num, world: int, rank: int, bias=None, flavour: int = -1):
        lo, hi = split_rows(weight.shape[0], world, rank)
        super().__init__(weight[lo:hi].contiguous(), reorder_index, select_num,
                         None if bias is None else bias[lo:hi].contiguous(), flavour)
        self.world, self.rank = world, rank


class RowParallelQLinear(QLinearShard):
    """K-split by channel ownership with the ARC residual columns kept on the owning shard; all-reduce of partials."""

    def __init__(self, weight, reorder_index, select_num, world: int, rank: int, bias=None, group=None, flavour: int = -1):
        lo, hi = split_rows(weight.shape[1], world, rank, 64)
        local_idx, ke = shard_reorder_index(reorder_index, select_num, world, rank)
        super().__init__(weight[:, lo:hi].contiguous(), local_idx, ke, bias if rank == 0 else None, flavour)
        self.world, self.rank, self.group = world, rank, group
        self.k_slice = (lo, hi)

    @torch.no_grad()
    def forward(self, x_local: torch.Tensor):
        """x_local: this rank's [M, KQ/world] slice of the activation.  Returns the all-reduced [M, N]."""
        y = super().forward(x_local)
        if self.world > 1:
            dist.all_reduce(y, op=dist.ReduceOp.SUM, group=self.group)
        return y
