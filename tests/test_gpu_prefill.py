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


@pytest.mark.parametrize("name,bound", [("qwen2.5-7b", 0.30), ("llama-3-8b", 0.30)])
def test_one_layer_parity_with_arc_columns(name, bound):
    """H = 3584 (qwen2.5-7b) is where the reference's KQ table disagrees with itself: rmsnorm_quantize_x is 16-wide
    (bindings.cpp:243), reorder_quantize_w 32-wide (:199).  With mismatched interleaves the ARC columns of X and W pair
    wrongly and the q/k/v/gate/up outputs are garbage (relative error ~1.0); the layer passes FLAVOUR_16 to the weights
    fed by rmsnorm_quantize_x, so one layer must sit within NVFP4 noise of the fp32 layer."""
    cfg = prefill.PrefillConfig.named(name, layers=1)
    dev = torch.device("cuda:0")
    model = prefill.ArcPrefillModel(cfg, dev, seed=5, keep_ref=True)
    g = torch.Generator(device=dev).manual_seed(1)
    h = torch.randn(1, 256, cfg.hidden, device=dev, generator=g).to(torch.bfloat16)
    y = model(h)
    ref = prefill.reference_layer_forward(model.layers[0], h)
    torch.cuda.synchronize()
    d, r = (y.float() - h.float()), (ref - h.float())
    rel = float((d - r).norm() / r.norm())
    cos = float(torch.nn.functional.cosine_similarity(d.flatten(), r.flatten(), dim=0))
    assert rel < bound and cos > 0.95, (name, rel, cos)


def test_h3584_qkv_outputs_are_not_garbage():
    """The column-parallel GEMM right after rmsnorm_quantize_x at H = 3584, KE = 256, against the fp32 product."""
    from arcquant_b200 import agemm, tp
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(2)
    H, KE, M, N = 3584, 256, 256, 512
    x = torch.randn(M, H, device=dev, generator=g).to(torch.bfloat16)
    x[:, :KE] *= 15
    idx = torch.argsort(x.float().abs().amax(0), stable=True).to(torch.int16)
    w = (torch.randn(N, H, device=dev, generator=g) * 0.02).to(torch.bfloat16)
    norm = torch.ones(H, dtype=torch.bfloat16, device=dev)
    qx, sfx = agemm.rmsnorm_quantize_x(x, norm, 1e-6, idx, KE)
    xn = x.float() * torch.rsqrt(x.float().pow(2).mean(-1, keepdim=True) + 1e-6)
    ref = xn @ w.float().T
    good = tp.QLinearShard(w, idx, KE, flavour=agemm.FLAVOUR_16).forward_quantized(qx, sfx, 1.0).float()
    bad = tp.QLinearShard(w, idx, KE, flavour=-1).forward_quantized(qx, sfx, 1.0).float()   # the reference's table: 32-wide
    torch.cuda.synchronize()
    rel_good, rel_bad = float((good - ref).norm() / ref.norm()), float((bad - ref).norm() / ref.norm())
    assert rel_good < 0.15, rel_good
    assert rel_bad > 3 * rel_good, "the table mismatch should be visible (%.3f vs %.3f)" % (rel_bad, rel_good)


def test_fused_launches_equal_the_reference_launch_structure():
    """4 + 4 launches per layer (fused q/k/v, SwiGLU epilogue, in-kernel pre-scale) against 4 + 7 launches with the
    torch glue of the reference's layers: same bytes at every step, so the same output."""
    from arcquant_b200 import agemm
    cfg = prefill.PrefillConfig.named("tiny")
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    h = torch.randn(2, 192, cfg.hidden, device=dev, generator=g).to(torch.bfloat16)
    fused = prefill.ArcPrefillModel(cfg, dev, seed=3, fused=True)
    unfused = prefill.ArcPrefillModel(cfg, dev, seed=3, fused=False)
    assert fused.layers[0].fused and not unfused.layers[0].fused
    n0 = agemm.launch_count()
    yf = fused(h)
    n1 = agemm.launch_count()
    yu = unfused(h)
    n2 = agemm.launch_count()
    torch.cuda.synchronize()
    assert n1 - n0 == cfg.layers * (4 + 2 + 4) and n2 - n1 == cfg.layers * (4 + 7)   # fused: + 2 absmax pre-passes
    diff = float((yf.float() - yu.float()).abs().max())
    assert torch.equal(yf, yu) or diff <= 2 ** -6 * float(yu.float().abs().max()), diff


def test_prefill_runs_on_calibration_artifacts(tmp_path):
    """The on-disk schema of the reference's calibration (./saved/*_{reorder_index,select_num}_*.pt) drives the layers."""
    from arcquant_b200 import calib as C
    cfg = prefill.PrefillConfig.named("tiny")
    ri, sn, sc = C.synthetic_artifacts(cfg.hidden, cfg.intermediate, cfg.layers, seed=1)
    C.save_artifacts("tiny", ri, sn, sc, root=str(tmp_path))
    ri, sn, _ = C.load_artifacts("tiny", root=str(tmp_path))
    dev = torch.device("cuda:0")
    model = prefill.ArcPrefillModel(cfg, dev, seed=3, keep_ref=True, calib=(ri, sn))
    assert model.layers[1].down.select_num == sn[C.key(1, "mlp", "down_proj")]
    assert torch.equal(model.layers[0].idx_h1.cpu(), ri[C.key(0, "self_attn", "q_proj")].to(torch.int16))
    g = torch.Generator(device=dev).manual_seed(0)
    h = torch.randn(2, 128, cfg.hidden, device=dev, generator=g).to(torch.bfloat16)
    y = model(h)
    ref = h
    for layer in model.layers:
        ref = prefill.reference_layer_forward(layer, ref)
    torch.cuda.synchronize()
    d, r = (y.float() - h.float()), (ref - h.float())
    assert float((d - r).norm() / r.norm()) < 0.35


def test_chained_forward_equals_layer_by_layer():
    """ArcPrefillModel.forward carries the down_proj output as a pending residual and adds it inside the NEXT norm's kernel
    (add_rmsnorm_quantize_x); calling the layers one by one does the adds in torch.  Same bits (the fused add is torch's
    bf16 add), for a model with q/k/v biases (bias epilogue) and one without."""
    dev = torch.device("cuda:0")
    for name in ("tiny", "llama-3-8b"):
        cfg = prefill.PrefillConfig.named(name, layers=2)
        model = prefill.ArcPrefillModel(cfg, dev, seed=7)
        g = torch.Generator(device=dev).manual_seed(2)
        h = torch.randn(2, 192, cfg.hidden, device=dev, generator=g).to(torch.bfloat16)
        y = model(h)
        z = h
        for layer in model.layers:
            z = layer(z)
        torch.cuda.synchronize()
        assert torch.equal(y, z), name
