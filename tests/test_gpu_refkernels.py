"""The reference's OWN quantise kernels (kernels/src/{reorder,down,rmsnorm}.cu compiled unchanged for sm_100a into
oracle/_ref/libarcq_refkernels.so by oracle/build_ref.sh) run on the B200 next to ours: bit-exact codes and scales.

This is the pin of the CPU oracle against the reference itself: reference kernel == oracle == ours.
Skipped when oracle/_ref was not built (the .so is git-ignored but travels with the gpurun snapshot).
"""
import ctypes
import os

import numpy as np
import pytest
import torch

import oracle
from tests._util import bits, make_acts, make_weights, sf_valid

pytestmark = pytest.mark.gpu

_SO = os.path.join(os.path.dirname(os.path.abspath(oracle.__file__)), "_ref", "libarcq_refkernels.so")


@pytest.fixture(scope="module")
def R():
    if not os.path.exists(_SO):
        pytest.skip("oracle/_ref/libarcq_refkernels.so not built (oracle/build_ref.sh kernels)")
    L = ctypes.CDLL(_SO)
    vp, i = ctypes.c_void_p, ctypes.c_int
    L.arcq_ref_quantize_x.argtypes = [vp, i, i, vp, i, vp, vp]
    L.arcq_ref_quantize_w.argtypes = [vp, i, i, vp, i, vp, vp]
    L.arcq_ref_rmsnorm_quantize_x.argtypes = [vp, vp, ctypes.c_float, i, i, vp, i, vp, vp]
    return L


def _ref_quant(R, fn, x, idx, KE):
    """Call the reference launcher the way bindings.cpp:122-163 does (index_select first for KQ in {27648, 28672})."""
    M, KQ = x.shape
    K = KQ + KE
    xd, idxd = x.cuda(), idx.cuda()
    if KQ in (27648, 28672):
        xd = torch.index_select(xd, 1, idxd.to(torch.int32)).contiguous()  # bindings.cpp:29-38
    q = torch.zeros((M, K // 2), dtype=torch.uint8, device="cuda")
    sf = torch.zeros(((M // 128 + 1) * 128 * K // 16,), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    rc = fn(xd.data_ptr(), M, KQ, idxd.data_ptr(), KE, q.data_ptr(), sf.data_ptr())
    assert rc == 0, "reference kernel failed: %d" % rc
    return q.cpu().numpy(), sf.cpu().numpy()


X_CASES = [(2500, 4096, 256, "A"), (1290, 14336, 512, "A"), (1300, 3584, 256, "A"), (1400, 8192, 512, "B"),  # multi-row tiles
           (128, 4096, 256, "A"), (300, 4096, 0, "A"), (77, 4096, 1024, "B"), (1, 4096, 64, "A"), (256, 2048, 128, "small"),
           (130, 14336, 512, "A"), (64, 3584, 256, "A"), (33, 18944, 512, "B"), (20, 28672, 1024, "A"), (40, 27648, 512, "A"),
           (129, 8192, 512, "A"), (50, 5120, 320, "A"), (16, 11008, 1024, "A"), (24, 13824, 512, "A"), (9, 3072, 192, "B")]


@pytest.mark.parametrize("M,KQ,KE,dist", X_CASES)
def test_x_ours_equals_reference_kernel_equals_oracle(R, M, KQ, KE, dist):
    from arcquant_b200 import agemm
    O = oracle.lib()
    x, idx = make_acts(M, KQ, KE, seed=M + KQ, dist=dist)
    q_ref, sf_ref = _ref_quant(R, R.arcq_ref_quantize_x, x, idx, KE)
    q, sf = agemm.reorder_quantize_x(x.cuda(), idx.cuda(), KE)
    torch.cuda.synchronize()
    q, sf = q.cpu().numpy(), sf.cpu().numpy()
    K = KQ + KE
    assert np.array_equal(q, q_ref), "%d code bytes differ from the reference kernel" % int((q != q_ref).sum())
    assert np.array_equal(sf_valid(O, sf, M, K), sf_valid(O, sf_ref, M, K))
    q_o, sf_o = O.quantize_x(bits(x), idx.numpy(), KE)
    assert np.array_equal(q_o, q_ref) and np.array_equal(sf_valid(O, sf_o, M, K), sf_valid(O, sf_ref, M, K)), \
        "CPU oracle differs from the reference kernel"


@pytest.mark.parametrize("N,KQ,KE", [(256, 4096, 256), (130, 3584, 256), (64, 14336, 512), (32, 28672, 1024),
                                     (40, 18944, 512), (128, 8192, 0)])
def test_w_ours_equals_reference_kernel_equals_oracle(R, N, KQ, KE):
    from arcquant_b200 import agemm
    O = oracle.lib()
    w, _ = make_weights(N, KQ, seed=N)
    _, idx = make_acts(8, KQ, KE, seed=3)
    q_ref, sf_ref = _ref_quant(R, R.arcq_ref_quantize_w, w, idx, KE)
    q, sf = agemm.reorder_quantize_w(w.cuda(), idx.cuda(), KE)
    torch.cuda.synchronize()
    q, sf = q.cpu().numpy(), sf.cpu().numpy()
    K = KQ + KE
    assert np.array_equal(q, q_ref)
    assert np.array_equal(sf_valid(O, sf, N, K), sf_valid(O, sf_ref, N, K))
    q_o, sf_o = O.quantize_w(bits(w), idx.numpy(), KE)
    assert np.array_equal(q_o, q_ref) and np.array_equal(sf_valid(O, sf_o, N, K), sf_valid(O, sf_ref, N, K))


@pytest.mark.parametrize("M,H,KE", [(128, 4096, 256), (37, 3584, 256), (200, 2048, 0), (64, 8192, 512), (9, 5120, 64),
                                    # prefill sizes: the shape-driven dispatch picks the multi-row tiles here (M >= 1184)
                                    (1300, 4096, 256), (2500, 3584, 256), (1290, 8192, 512), (1184, 2048, 128), (1500, 5120, 320),
                                    (4096 + 77, 4096, 256)])
def test_rmsnorm_ours_equals_reference_kernel(R, M, H, KE):
    """Same device, same rsqrtf, same reduction order -> the fused rmsnorm variant must match bit for bit."""
    from arcquant_b200 import agemm
    O = oracle.lib()
    g = torch.Generator().manual_seed(H + M)
    x = (torch.randn(M, H, generator=g) * 3).to(torch.bfloat16)
    x[:, :KE] *= 10
    w = (1.0 + 0.1 * torch.randn(H, generator=g)).to(torch.bfloat16)
    idx = torch.argsort(x.float().abs().amax(0), stable=True).to(torch.int16)
    K = H + KE
    xd, wd, idxd = x.cuda(), w.cuda(), idx.cuda()
    q_ref = torch.zeros((M, K // 2), dtype=torch.uint8, device="cuda")
    sf_ref = torch.zeros(((M // 128 + 1) * 128 * K // 16,), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    rc = R.arcq_ref_rmsnorm_quantize_x(xd.data_ptr(), wd.data_ptr(), 1e-6, M, H, idxd.data_ptr(), KE, q_ref.data_ptr(),
                                       sf_ref.data_ptr())
    assert rc == 0
    q, sf = agemm.rmsnorm_quantize_x(xd, wd, 1e-6, idxd, KE)
    torch.cuda.synchronize()
    assert torch.equal(q, q_ref)
    assert np.array_equal(sf_valid(O, sf.cpu().numpy(), M, K), sf_valid(O, sf_ref.cpu().numpy(), M, K))
