"""Prefill harness (arcquant_b200/prefill.py) against the same decoder layers in fp32 torch with unquantised weights:
NVFP4+ARC is a lossy format, so the bound is statistical (relative Frobenius error and cosine), not bitwise."""
import pytest
import torch

from arcquant_b200 import prefill

pytestmark = pytest.mark.gpu


def test_tiny_prefill_matches_fp32_reference():
    cfg = prefill.PrefillConfig.named("tiny")
    dev = torch.device("cuda:0")
    model = prefill.ArcPrefillModel(cfg, dev, seed=3, keep_ref=True)
    g = torch.Generator(device=dev).manual_seed(0)
    h = torch.randn(2, 128, cfg.hidden, device=dev, generator=g).to(torch.bfloat16)
    y = model(h)
    torch.cuda.synchronize()
    ref = h
    for layer in model.layers:
        ref = prefill.reference_layer_forward(layer, ref)
    assert y.shape == h.shape and y.dtype == torch.bfloat16 and torch.isfinite(y).all()
    d, r = (y.float() - h.float()), (ref - h.float())  # compare what the layers ADD to the residual stream
    rel = (d - r).norm() / r.norm()
    cos = torch.nn.functional.cosine_similarity(d.flatten(), r.flatten(), dim=0)
    # 4-bit operands on both sides give ~13 % error per GEMM; seven chained linears per layer, two layers (measured 0.26 / 0.965)
    assert rel < 0.35 and cos > 0.94, (float(rel), float(cos))


def test_prefill_shapes_of_the_named_models():
    for name in ("qwen2.5-7b", "llama-3-8b"):
        cfg = prefill.PrefillConfig.named(name, layers=1)
        model = prefill.ArcPrefillModel(cfg, torch.device("cuda:0"))
        h = torch.randn(1, 256, cfg.hidden, device="cuda").to(torch.bfloat16)
        y = model(h)
        torch.cuda.synchronize()
        assert y.shape == h.shape and torch.isfinite(y).all()
        assert model.linear_flops(256) > 0
