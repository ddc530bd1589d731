"""Every tile organisation of the TMA-staged quantise kernel against the oracle (GPU).

The shape-driven dispatch (quant_tma.cu: launch_quant_tma) picks the 1-row tile for the small M the other parity tests
use, so the 4- and 2-row tiles that real prefill sizes run (ldmatrix/stmatrix transposition, owned gather offsets,
packed two-row arithmetic, the 64-channel tail step) are forced here through the experiment knob
arcq_set_option("quant_tile", "R,threads,S,RT,CTAs/SM"), and the default dispatch is checked once at a size where it picks them itself.
Bit-exact against oracle/ (reference kernels: kernels/src/reorder.cu, down.cu, rmsnorm.cu).
"""
import numpy as np
import pytest
import torch

from tests._util import bits, make_acts
import oracle

pytestmark = pytest.mark.gpu


def _check(q, sf, q_ref, sf_ref, M, K, what):
    off = oracle.lib().sf_valid_offsets(M, K)
    a, b = sf.cpu().numpy()[off], sf_ref[off]
    nsf = int((a != b).sum())
    nq = int((q.cpu().numpy() != q_ref).sum())
    assert nsf == 0 and nq == 0, "%s: %d scale bytes, %d code bytes differ" % (what, nsf, nq)


TILE_CASES = [
    # knob,            M,    KQ,    KE,  flavour
    ("4,256,2,4,2", 1300, 4096, 256, None),   # the prefill default: 4-row tiles, two CTAs per SM, ragged last block
    ("4,256,1,4,2", 515, 4096, 256, None),    # single staging slot
    ("4,256,2,4,2", 260, 3584, 256, None),    # 32-wide flavour, KQ/16 < threads
    ("4,256,2,4,2", 140, 2048, 128, None),
    ("4,512,2,4,1", 700, 8192, 512, None),    # 4-row tiles, one wide CTA per SM
    ("4,512,2,4,1", 129, 4096, 4096, None),   # every channel has a residual
    ("2,256,2,2,2", 333, 9280, 320, 0),       # 2-row tiles, KQ = 64 (mod 128): the two-matrix tail step
    ("2,256,2,2,2", 200, 4096, 256, None),
    ("2,512,2,2,1", 600, 14336, 512, None),   # gather offsets from the shared table (more groups than threads)
    ("2,512,2,2,1", 300, 18944, 512, None),   # 32-wide flavour
    ("2,512,1,2,1", 130, 18944, 512, None),
    ("1,512,1,1,1", 100, 28672, 1024, None),
    ("8,512,1,8,1", 260, 4096, 256, None),    # 8-row tiles (not chosen by default)
    ("8,512,2,4,1", 300, 4096, 256, None),    # 8-row tiles, two threads per group
    ("8,512,1,4,1", 1290, 4096, 256, None),   # the default for KQ <= 4096 at prefill sizes
    ("8,512,1,4,1", 200, 3584, 256, None),
    ("8,512,1,4,1", 150, 2048, 0, None),
]


@pytest.mark.parametrize("knob,M,KQ,KE,flavour", TILE_CASES)
def test_forced_tile_quantize_x(arcq_option, knob, M, KQ, KE, flavour):
    from arcquant_b200 import agemm
    O = oracle.lib()
    x, idx = make_acts(M, KQ, KE, seed=M + KQ, dist="A")
    q_ref, sf_ref = O.quantize_x(bits(x), idx.numpy(), KE, flavour)
    arcq_option("quant_tile", knob)
    q, sf = agemm.reorder_quantize_x(x.cuda(), idx.cuda(), KE, -1 if flavour is None else flavour)
    torch.cuda.synchronize()
    _check(q, sf, q_ref, sf_ref, M, KQ + KE, "quantize_x %s M=%d KQ=%d KE=%d" % (knob, M, KQ, KE))


@pytest.mark.parametrize("knob,N,KQ,KE", [("4,256,2,4,2", 520, 4096, 256), ("8,512,1,4,1", 300, 4096, 256), ("2,512,2,2,1", 260, 14336, 512),
                                          ("4,512,2,4,1", 300, 8192, 512)])
def test_forced_tile_quantize_w(arcq_option, knob, N, KQ, KE):
    from arcquant_b200 import agemm
    O = oracle.lib()
    w, idx = make_acts(N, KQ, KE, seed=N + 1, dist="B")
    q_ref, sf_ref = O.quantize_w(bits(w), idx.numpy(), KE, None)
    arcq_option("quant_tile", knob)
    q, sf = agemm.reorder_quantize_w(w.cuda(), idx.cuda(), KE)
    torch.cuda.synchronize()
    _check(q, sf, q_ref, sf_ref, N, KQ + KE, "quantize_w %s N=%d KQ=%d KE=%d" % (knob, N, KQ, KE))


@pytest.mark.parametrize("M,KQ,KE", [(2500, 4096, 256), (1290, 14336, 512), (1400, 8192, 512), (1300, 3584, 256)])
def test_default_dispatch_at_prefill_sizes(M, KQ, KE):
    """No knob: M large enough that the dispatch itself picks the multi-row tiles."""
    from arcquant_b200 import agemm
    O = oracle.lib()
    x, idx = make_acts(M, KQ, KE, seed=KQ, dist="A")
    q_ref, sf_ref = O.quantize_x(bits(x), idx.numpy(), KE, None)
    q, sf = agemm.reorder_quantize_x(x.cuda(), idx.cuda(), KE)
    torch.cuda.synchronize()
    _check(q, sf, q_ref, sf_ref, M, KQ + KE, "quantize_x default M=%d KQ=%d KE=%d" % (M, KQ, KE))


RMS_CASES = [("8,512,1,4,1", 1300, 4096, 256), ("8,512,1,4,1", 140, 3584, 256), ("4,256,1,4,2", 1300, 4096, 256), ("4,256,1,4,2", 200, 3584, 256), ("4,512,1,4,1", 300, 8192, 512),
             ("2,512,2,2,1", 300, 8192, 512), ("4,256,2,4,2", 260, 2048, 128), ("2,256,2,2,2", 150, 5120, 64)]


@pytest.mark.parametrize("knob,M,H,KE", RMS_CASES)
def test_forced_tile_rmsnorm_matches_one_row_tiles(arcq_option, knob, M, H, KE):
    """The 1-row tile kernel is pinned to the oracle by test_gpu_quant.py (with its rsqrt-ulp allowance); every other
    tile height runs the same reduction tree, so it must reproduce the 1-row result bit for bit."""
    from arcquant_b200 import agemm
    g = torch.Generator().manual_seed(H + M)
    x = (torch.randn(M, H, generator=g) * 3).to(torch.bfloat16).cuda()
    w = (1.0 + 0.1 * torch.randn(H, generator=g)).to(torch.bfloat16).cuda()
    idx = torch.randperm(H, generator=g).to(torch.int16).cuda()
    arcq_option("quant_tile", "1,512,1,1,1")
    q1, sf1 = agemm.rmsnorm_quantize_x(x, w, 1e-6, idx, KE)
    arcq_option("quant_tile", knob)
    q2, sf2 = agemm.rmsnorm_quantize_x(x, w, 1e-6, idx, KE)
    torch.cuda.synchronize()
    off = oracle.lib().sf_valid_offsets(M, H + KE)
    assert torch.equal(q1, q2), "%s: codes differ from the 1-row tile kernel" % knob
    assert np.array_equal(sf1.cpu().numpy()[off], sf2.cpu().numpy()[off]), "%s: scales differ" % knob


def test_default_rmsnorm_at_prefill_size_matches_one_row_tiles(arcq_option):
    from arcquant_b200 import agemm
    M, H, KE = 2500, 4096, 256
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(M, H, generator=g) * 3).to(torch.bfloat16).cuda()
    w = (1.0 + 0.1 * torch.randn(H, generator=g)).to(torch.bfloat16).cuda()
    idx = torch.randperm(H, generator=g).to(torch.int16).cuda()
    q2, sf2 = agemm.rmsnorm_quantize_x(x, w, 1e-6, idx, KE)
    arcq_option("quant_tile", "1,512,1,1,1")
    q1, sf1 = agemm.rmsnorm_quantize_x(x, w, 1e-6, idx, KE)
    torch.cuda.synchronize()
    off = oracle.lib().sf_valid_offsets(M, H + KE)
    assert torch.equal(q1, q2)
    assert np.array_equal(sf1.cpu().numpy()[off], sf2.cpu().numpy()[off])


@pytest.mark.parametrize("rms", [False, True])
def test_repeatable_under_load(rms):
    """Race hunt by repetition (compute-sanitizer is not available on the pool): back-to-back launches over alternating
    inputs must reproduce their first result bit for bit -- a staging slot refilled before its last reader, or a scale
    buffer copied out a tile early, shows up as a sporadic difference."""
    from arcquant_b200 import agemm
    M, KQ, KE = 4096 + 77, 4096, 256
    g = torch.Generator().manual_seed(11)
    xs = [(torch.randn(M, KQ, generator=g) * (1 + 3 * i)).to(torch.bfloat16).cuda() for i in range(3)]
    idx = torch.randperm(KQ, generator=g).to(torch.int16).cuda()
    w = (1.0 + 0.1 * torch.randn(KQ, generator=g)).to(torch.bfloat16).cuda()

    def run(x):
        return agemm.rmsnorm_quantize_x(x, w, 1e-6, idx, KE) if rms else agemm.reorder_quantize_x(x, idx, KE)

    first = [tuple(t.clone() for t in run(x)) for x in xs]
    off = torch.from_numpy(oracle.lib().sf_valid_offsets(M, KQ + KE).astype(np.int64)).cuda()
    outs = [run(xs[i % 3]) for i in range(45)]
    torch.cuda.synchronize()
    for i, (q, sf) in enumerate(outs):
        assert torch.equal(q, first[i % 3][0]), "codes of launch %d differ from the first run" % i
        assert torch.equal(sf[off], first[i % 3][1][off]), "scales of launch %d differ from the first run" % i
