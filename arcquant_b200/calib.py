"""Calibration artefacts of the reference, as the hot path consumes them (SURVEY.md §5 "Checkpoint / resume", §8 f4).

The reference's calibration (reorder_indices.py / utilize.py -- OUT OF SCOPE here: it needs checkpoints and datasets)
leaves three ``torch.save``d dicts under ``./saved/`` that its model wrappers load at start-up:

    {model}_reorder_index_{dataset}_{metric}.pt   name -> int64 permutation, ascending activation scale so the
                                                   outliers come LAST (utilize.py:42, reorder_indices.py:88-93)
    {model}_select_num_{dataset}_{metric}.pt      name -> int, a multiple of 64 (utilize.py:475)
    {model}_act_scales_{dataset}_{metric}.pt      name -> per-channel statistic (reorder_indices.py:68)

with ``name = "layers.{i}.{self_attn|mlp}.{proj}.input"`` (utilize.py:53, model/qLlamaLayer.py:237,
benchmarks/modeling_arc.py:95-106,346-347).  This module reads and writes exactly that schema, validates it, and can
synthesise it (no network here), so that reference callers -- and this repo's prefill harness -- run unmodified on real or
synthetic calibration.  Pure host-side bookkeeping: nothing here touches a kernel.
"""
from __future__ import annotations

import math
import os
from typing import Dict, Optional, Tuple

import torch

NAME = "layers.{}.{}.{}.{}"  # the reference's nameTemplate (model/model_utils.py:38, qLlamaLayer.py:237)
ATTN_PROJS = ("q_proj", "k_proj", "v_proj", "o_proj")
MLP_PROJS = ("gate_proj", "up_proj", "down_proj")
SELECT_ALIGN = 64  # utilize.py:475


def key(layer: int, block: str, proj: str) -> str:
    return NAME.format(layer, block, proj, "input")


def artifact_paths(model_name: str, dataset: str = "wikitext2", metric: str = "max", root: str = "./saved") -> Dict[str, str]:
    """File names of reorder_indices.py:88-90 / model/main.py:112-114."""
    stem = os.path.join(root, model_name.lower())
    return {k: "%s_%s_%s_%s.pt" % (stem, k, dataset.lower(), metric) for k in ("reorder_index", "select_num", "act_scales", "average_bits")}


def select_num_for(in_features: int, ratio: float) -> int:
    """select_num = ceil(in_features * ratio / 64) * 64 (utilize.py:475)."""
    return int(math.ceil(in_features * ratio / SELECT_ALIGN) * SELECT_ALIGN)


def reorder_from_scales(act_scale: torch.Tensor) -> torch.Tensor:
    """Ascending sort of the per-channel statistic: outliers last (utilize.py:42)."""
    return torch.sort(act_scale.float(), descending=False, stable=True)[1]


def validate(reorder_index: Dict[str, torch.Tensor], select_nums: Dict[str, int]) -> None:
    """Raise ValueError unless every entry is usable by the kernels: a permutation of [0, KQ) that fits int16
    (model_utils.py:39-42 casts to int16), select_num a multiple of 64 with 0 <= select_num <= KQ."""
    for name, idx in reorder_index.items():
        kq = idx.numel()
        if idx.dim() != 1 or kq > 32768:
            raise ValueError("%s: reorder index must be 1-D with at most 32768 entries" % name)
        if not torch.equal(torch.sort(idx.to(torch.int64))[0], torch.arange(kq)):
            raise ValueError("%s: not a permutation of [0, %d)" % (name, kq))
        if name not in select_nums:
            raise ValueError("%s: no select_num" % name)
        ke = int(select_nums[name])
        if ke % SELECT_ALIGN or not 0 <= ke <= kq:
            raise ValueError("%s: select_num=%d must be a multiple of %d in [0, %d]" % (name, ke, SELECT_ALIGN, kq))


def load_artifacts(model_name: str, dataset: str = "wikitext2", metric: str = "max", root: str = "./saved",
                   check: bool = True) -> Tuple[Dict[str, torch.Tensor], Dict[str, int], Optional[Dict[str, torch.Tensor]]]:
    """(reorder_index, select_nums, act_scales or None) as model/main.py:120-122 loads them."""
    paths = artifact_paths(model_name, dataset, metric, root)
    for k in ("reorder_index", "select_num"):
        if not os.path.isfile(paths[k]):
            raise FileNotFoundError("%s file not found: %s" % (k.replace("_", " "), paths[k]))  # model/main.py:117
    reorder_index = torch.load(paths["reorder_index"], map_location="cpu", weights_only=True)
    select_nums = {k: int(v) for k, v in torch.load(paths["select_num"], map_location="cpu", weights_only=True).items()}
    act_scales = torch.load(paths["act_scales"], map_location="cpu", weights_only=True) if os.path.isfile(paths["act_scales"]) else None
    if check:
        validate(reorder_index, select_nums)
    return reorder_index, select_nums, act_scales


def save_artifacts(model_name: str, reorder_index, select_nums, act_scales=None, dataset: str = "wikitext2", metric: str = "max",
                   root: str = "./saved") -> Dict[str, str]:
    """Write the files reorder_indices.py:88-97 writes (average_bits included: 4.5 * (in + select) / in, utilize.py:477)."""
    os.makedirs(root, exist_ok=True)
    paths = artifact_paths(model_name, dataset, metric, root)
    torch.save({k: v.to(torch.int64) for k, v in reorder_index.items()}, paths["reorder_index"])
    torch.save({k: int(v) for k, v in select_nums.items()}, paths["select_num"])
    torch.save({k: 4.5 * (reorder_index[k].numel() + int(v)) / reorder_index[k].numel() for k, v in select_nums.items()}, paths["average_bits"])
    if act_scales is not None:
        torch.save(act_scales, paths["act_scales"])
    return paths


def synthetic_artifacts(hidden: int, intermediate: int, n_layers: int, q_out: Optional[int] = None, ratio_hidden: float = 1 / 16,
                        ratio_intermediate: float = 1 / 28, seed: int = 0, outlier_gain: float = 20.0):
    """The three dicts for a Llama/Qwen-shaped stack with synthetic statistics: per-channel scales ~ |N(0,1)| with a
    heavy-tailed minority, index = ascending sort, select_num by the reference's rounding rule.  q/k/v share the statistic
    of their common input, gate/up likewise (they are hooks on the same tensor in the reference, utilize.py:49-53)."""
    g = torch.Generator().manual_seed(seed)
    q_out = hidden if q_out is None else q_out

    def stat(n, ratio):
        s = torch.randn(n, generator=g).abs() + 0.05
        hot = torch.randperm(n, generator=g)[: max(1, int(n * ratio))]
        s[hot] *= outlier_gain
        return s

    act_scales, reorder_index, select_nums = {}, {}, {}
    for i in range(n_layers):
        shared = {"attn_in": stat(hidden, ratio_hidden), "o_in": stat(q_out, ratio_hidden), "mlp_in": stat(hidden, ratio_hidden),
                  "down_in": stat(intermediate, ratio_intermediate)}
        plan = [("self_attn", "q_proj", "attn_in", ratio_hidden), ("self_attn", "k_proj", "attn_in", ratio_hidden),
                ("self_attn", "v_proj", "attn_in", ratio_hidden), ("self_attn", "o_proj", "o_in", ratio_hidden),
                ("mlp", "gate_proj", "mlp_in", ratio_hidden), ("mlp", "up_proj", "mlp_in", ratio_hidden),
                ("mlp", "down_proj", "down_in", ratio_intermediate)]
        for block, proj, src, ratio in plan:
            k = key(i, block, proj)
            act_scales[k] = shared[src]
            reorder_index[k] = reorder_from_scales(shared[src])
            select_nums[k] = select_num_for(shared[src].numel(), ratio)
    return reorder_index, select_nums, act_scales


def layer_entries(reorder_index, select_nums, layer: int, device=None):
    """The four quantise configurations of one decoder layer, as the reference's layers use them
    (model_utils.py:39-42, qLlamaLayer.py:296-297,364,429,439): name -> (int16 index on `device`, select_num).
    q/k/v share q_proj's entry, gate/up share up_proj's."""
    out = {}
    for tag, block, proj in (("qkv", "self_attn", "q_proj"), ("o", "self_attn", "o_proj"), ("gate_up", "mlp", "up_proj"),
                             ("down", "mlp", "down_proj")):
        k = key(layer, block, proj)
        idx = reorder_index[k].to(torch.int16)
        out[tag] = (idx.to(device) if device is not None else idx, int(select_nums[k]))
    return out
