"""Fused per-tensor pre-scale: `arcq_absmax` + `arcq_quantize_x_scaled` must reproduce, byte for byte,
what the reference callers do in torch in front of the kernel (model/qLlamaLayer.py:73-77, benchmarks/modeling_arc.py:58-61):

    scale = torch.max(x.abs()).float() / (448.0 * 6.0);  qx, sfx = agemm.reorder_quantize_x(x / scale, idx, KE)
"""
import numpy as np
import pytest
import torch

import oracle
from tests._util import bits

pytestmark = pytest.mark.gpu


def _raw_acts(M, KQ, KE, seed, mag=1.0):
    """UN-scaled activations (the callers' input) with heavy-tailed outlier channels + the ascending-absmax order."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, KQ, generator=g) * mag
    perm = torch.randperm(KQ, generator=g)
    if KE:
        x[:, perm[:KE]] *= 20.0
    x[::7, ::5] = 0.0
    x[1::9, 3::11] = -0.0
    x = x.to(torch.bfloat16)
    idx = torch.argsort(x.float().abs().amax(0), stable=True).to(torch.int16)
    return x.contiguous(), idx.contiguous()


def _torch_path(agemm, x, idx, KE, flavour=-1, literal=False):
    """literal=True: the reference's expression verbatim (0-dim CUDA divisor, qLlamaLayer.py:74-76); False: the same
    division with the divisor kept in fp32 (what this repo's own callers do: arcquant_b200/tp.py)."""
    scale = torch.max(x.abs()).float() / (448.0 * 6.0)
    xs = x / scale if literal else (x / scale.reshape(1)).to(torch.bfloat16)
    assert xs.dtype == torch.bfloat16
    q, sf = agemm.reorder_quantize_x(xs, idx, KE, flavour)
    return q, sf, scale


def test_which_division_torch_does_for_a_0dim_cuda_divisor():
    """Documents the semantics ARCQ_PRESCALE_BF16_DIVISOR exists for: on CUDA, `bf16_tensor / fp32_0dim_tensor` casts the
    divisor to bf16 first (TensorIterator common dtype); a 1-element 1-D divisor promotes the result to fp32 instead."""
    g = torch.Generator(device="cuda").manual_seed(0)
    x = (torch.randn(1 << 16, device="cuda", generator=g) * 5).to(torch.bfloat16)
    scale = torch.tensor(0.0123456789, dtype=torch.float32, device="cuda")
    y = x / scale
    fp32_div = (x.float() / scale).to(torch.bfloat16)
    bf16_div = (x.float() / scale.to(torch.bfloat16).float()).to(torch.bfloat16)
    assert y.dtype == torch.bfloat16
    assert torch.equal(y, bf16_div) != torch.equal(y, fp32_div)
    assert torch.equal(y, bf16_div), "torch divides by the fp32 value here: make PRESCALE_FP32 the mode of the literal expression"
    assert torch.equal((x / scale.reshape(1)).to(torch.bfloat16), fp32_div)


@pytest.mark.parametrize("n", [1, 7, 8, 4096 * 3 + 5, 8192 * 4096])
def test_absmax_equals_torch(n):
    from arcquant_b200 import agemm
    g = torch.Generator(device="cuda").manual_seed(n)
    x = (torch.randn(n, device="cuda", generator=g) * 3).to(torch.bfloat16)
    if n > 8:
        x[n // 2] = -77.5
    got = agemm.absmax(x)
    assert got.dtype == torch.float32 and got.shape == (1,)
    assert float(got) == float(x.abs().max().float())
    assert float(agemm.absmax(torch.zeros(64, dtype=torch.bfloat16, device="cuda"))) == 0.0


SCALED_CASES = [(300, 4096, 256, 1.0), (2500, 4096, 256, 0.02), (1290, 14336, 512, 3.0), (130, 3584, 256, 1.0), (1300, 8192, 512, 50.0),
                (64, 2048, 0, 1e-3), (129, 18944, 512, 1.0), (1, 4096, 64, 1.0), (700, 5120, 320, 7.0), (200, 4096, 4096, 1.0)]


@pytest.mark.parametrize("M,KQ,KE,mag", SCALED_CASES)
def test_scaled_quantise_equals_torch_prescale_then_quantise(M, KQ, KE, mag):
    from arcquant_b200 import agemm
    O = oracle.lib()
    x, idx = _raw_acts(M, KQ, KE, seed=M + KQ, mag=mag)
    xd, idxd = x.cuda(), idx.cuda()
    off = torch.from_numpy(O.sf_valid_offsets(M, KQ + KE)).cuda()
    for literal, mode in ((False, agemm.PRESCALE_FP32), (True, agemm.PRESCALE_BF16_DIVISOR)):
        q_ref, sf_ref, scale_ref = _torch_path(agemm, xd, idxd, KE, literal=literal)
        q, sf, scale = agemm.reorder_quantize_x_scaled(xd, idxd, KE, mode=mode)
        torch.cuda.synchronize()
        assert float(scale) == float(scale_ref)
        nq, ns = int((q != q_ref).sum()), int((sf[off] != sf_ref[off]).sum())
        assert nq == 0 and ns == 0, "M=%d KQ=%d KE=%d mode=%d: %d code bytes, %d scale bytes differ" % (M, KQ, KE, mode, nq, ns)
    q, sf, scale = agemm.reorder_quantize_x_scaled(xd, idxd, KE, mode=agemm.PRESCALE_FP32)
    # and the whole chain against the CPU oracle fed with torch's (CPU) pre-scaled tensor
    # CUDA torch computes `tensor / python_scalar` as tensor * (1 / scalar) in fp32 (CPU torch divides): restate that
    s_cpu = torch.max(x.abs()).float() * torch.tensor(1.0, dtype=torch.float32).div(2688.0)
    assert float(s_cpu) == float(scale)
    q_o, sf_o = O.quantize_x(bits((x.float() / s_cpu).to(torch.bfloat16)), idx.numpy(), KE)
    offn = O.sf_valid_offsets(M, KQ + KE)
    assert np.array_equal(q.cpu().numpy(), q_o) and np.array_equal(sf.cpu().numpy()[offn], sf_o[offn])


def test_prescale_division_is_exact_for_every_bf16_input():
    """x / scale for ALL 65536 bf16 bit patterns x many scales: the in-kernel reciprocal + two-FMA division (and its slow
    path for all-ones significands / extreme magnitudes) must give torch's fp32-division-then-bf16 result wherever the
    result can influence a code or a scale, i.e. the quantised bytes must be identical."""
    from arcquant_b200 import agemm
    O = oracle.lib()
    KQ = 4096
    allbits = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)
    finite = allbits[torch.isfinite(allbits.float())]
    g = torch.Generator().manual_seed(1)
    reps = (16 * KQ + finite.numel() - 1) // finite.numel()
    base = finite.repeat(reps)[torch.randperm(reps * finite.numel(), generator=g)][: 16 * KQ].view(16, KQ).clone()
    idx = torch.randperm(KQ, generator=g).to(torch.int16).cuda()
    off = torch.from_numpy(O.sf_valid_offsets(16, KQ + 256)).cuda()
    # tensor maxima whose scale = amax / 2688 has an all-ones significand (the Markstein exception), extreme ones, ordinary ones
    amaxes = [2688.0 * float(np.float32(np.uint32(0x3f7fffff).view(np.float32))), 1.0, 3.0, 517.0, 1e-3, 2.0 ** -45, 2.0 ** 70, 12345.678, 0.3337]
    for amax in amaxes:
        x = base.clone().float()
        x = torch.where(x.abs() > amax, torch.sign(x) * amax * 0.5, x)   # keep |x| <= amax
        x[0, 0] = amax
        x = x.to(torch.bfloat16).cuda()
        for literal, mode in ((False, agemm.PRESCALE_FP32), (True, agemm.PRESCALE_BF16_DIVISOR)):
            q_ref, sf_ref, s_ref = _torch_path(agemm, x, idx, 256, literal=literal)
            q, sf, s = agemm.reorder_quantize_x_scaled(x, idx, 256, mode=mode)
            torch.cuda.synchronize()
            assert float(s) == float(s_ref)
            assert torch.equal(q, q_ref) and torch.equal(sf[off], sf_ref[off]), \
                "amax=%g mode=%d: %d code bytes differ" % (amax, mode, int((q != q_ref).sum()))


def test_scaled_quantise_accepts_a_carried_amax():
    """amax may come from the producer (e.g. a previous kernel): passing it skips the pre-pass."""
    from arcquant_b200 import agemm
    x, idx = _raw_acts(256, 4096, 256, seed=9)
    xd, idxd = x.cuda(), idx.cuda()
    am = agemm.absmax(xd)
    n0 = agemm.launch_count()
    q1, sf1, s1 = agemm.reorder_quantize_x_scaled(xd, idxd, 256, amax=am)
    assert agemm.launch_count() - n0 == 1
    q2, sf2, s2 = agemm.reorder_quantize_x_scaled(xd, idxd, 256)
    torch.cuda.synchronize()
    assert torch.equal(q1, q2) and float(s1) == float(s2)
