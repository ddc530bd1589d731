"""tcgen05 NVFP4 GEMM (through the C ABI) against the oracle's float64 dequantise-then-matmul.

Tolerance: the GPU accumulates in fp32 in hardware order and rounds once to bf16, so
|y - ref| <= 2^-8 |ref| (bf16 half-ulp is 2^-9; 2^-8 leaves room for accumulation-order noise) plus an
absolute floor of 2^-8 * alpha * sum_k |a_k b_k| * 2^-10 for cancelling sums.
"""
import numpy as np
import pytest
import torch

import oracle
from tests._util import bits, make_acts, make_weights

pytestmark = pytest.mark.gpu


def _ref_and_bound(O, qa, sfa, qb, sfb, K, alpha):
    da = O.dequant(qa, sfa, K).astype(np.float64)
    db = O.dequant(qb, sfb, K).astype(np.float64)
    ref = alpha * (da @ db.T)
    mag = alpha * (np.abs(da) @ np.abs(db).T)
    return ref, mag


def _run_case(M, N, KQ, KE, seed=0):
    from arcquant_b200 import agemm
    O = oracle.lib()
    x, idx = make_acts(M, KQ, KE, seed=seed)
    w, sw = make_weights(N, KQ, seed=seed + 1)
    qa, sfa = agemm.reorder_quantize_x(x.cuda(), idx.cuda(), KE)
    qb, sfb = agemm.reorder_quantize_w(w.cuda(), idx.cuda(), KE)
    alpha = 0.37 * sw
    y = agemm.matmul(qa, qb, sfa, sfb, alpha)
    torch.cuda.synchronize()
    assert y.shape == (M, N) and y.dtype == torch.bfloat16
    K = KQ + KE
    ref, mag = _ref_and_bound(O, qa.cpu().numpy(), sfa.cpu().numpy(), qb.cpu().numpy(), sfb.cpu().numpy(), K,
                              float(np.float32(alpha)))
    got = y.float().cpu().numpy().astype(np.float64)
    err = np.abs(got - ref)
    tol = 2.0 ** -8 * np.abs(ref) + 2.0 ** -18 * mag + 1e-30
    bad = err > tol
    rel = np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-30)
    assert not bad.any(), "M=%d N=%d K=%d: %d of %d outside tolerance, max err/tol %.3g, rel-F %.3g" % (
        M, N, K, int(bad.sum()), bad.size, float((err / tol).max()), rel)
    assert rel < 2e-3


@pytest.mark.parametrize("M,N,KQ,KE", [
    (128, 128, 256, 0),       # one tile, one k-block
    (128, 256, 256, 64),      # K tail (320 = 256 + 64)
    (128, 128, 1024, 0),
    (200, 384, 1024, 64),     # ragged M and N
    (1, 128, 1024, 128),
    (77, 1000, 2048, 128),    # N not a multiple of 8/128
    (512, 1024, 4096, 256),
    (300, 4096, 4096, 64),
    (256, 512, 14336, 512),
])
def test_gemm_vs_dequant_matmul(M, N, KQ, KE):
    _run_case(M, N, KQ, KE, seed=M + N)


VARIANTS = {"pair": {"gemm_pair": 1, "gemm_splitk": 0}, "bn128": {"gemm_pair": 0, "gemm_bn": 128, "gemm_splitk": 0},
            "bn256": {"gemm_pair": 0, "gemm_bn": 256, "gemm_splitk": 0}}


@pytest.mark.parametrize("variant", sorted(VARIANTS))
@pytest.mark.parametrize("M,N,KQ,KE", [(256, 256, 256, 0), (300, 520, 1024, 64), (77, 1000, 2048, 128), (1024, 768, 4096, 256)])
def test_gemm_kernel_variants(arcq_option, variant, M, N, KQ, KE):
    """Every kernel family (cta_group::2 pair, single-CTA 128/256-wide) on ragged shapes."""
    for k, v in VARIANTS[variant].items():
        arcq_option(k, v)
    _run_case(M, N, KQ, KE, seed=M + N + 1)


@pytest.mark.parametrize("bn", [128, 256])
def test_gemm_device_alpha_and_identity(bn):
    """Known answer: codes 1.0 (0x2) everywhere, scales 1.0 (0x38): y = alpha * K."""
    from arcquant_b200 import agemm
    M, N, K = 256, 512, 512
    a = torch.full((M, K // 2), 0x22, dtype=torch.uint8, device="cuda")
    b = torch.full((N, K // 2), 0x22, dtype=torch.uint8, device="cuda")
    sfa = torch.full((agemm.sf_bytes(M, K),), 0x38, dtype=torch.uint8, device="cuda")
    sfb = torch.full((agemm.sf_bytes(N, K),), 0x38, dtype=torch.uint8, device="cuda")
    alpha = torch.tensor([0.5], dtype=torch.float32, device="cuda")
    y = agemm.matmul(a, b, sfa, sfb, alpha)
    torch.cuda.synchronize()
    assert torch.all(y.float() == 0.5 * K), y


def test_gemm_linearity_full_size():
    """Full Llama-3-8B shape, size-independent property: halving every A scale and doubling alpha leaves y unchanged."""
    from arcquant_b200 import agemm
    M, N, KQ, KE = 2048, 4096, 4096, 256
    x, idx = make_acts(M, KQ, KE, seed=5)
    w, sw = make_weights(N, KQ, seed=6)
    qa, sfa = agemm.reorder_quantize_x(x.cuda(), idx.cuda(), KE)
    qb, sfb = agemm.reorder_quantize_w(w.cuda(), idx.cuda(), KE)
    y1 = agemm.matmul(qa, qb, sfa, sfb, 1e-4)
    valid = (M + 127) // 128 * 128 * (KQ + KE) // 16
    if bool((sfa[:valid] < 0x10).any()):
        pytest.skip("scales too small to halve exactly")
    sfa2 = sfa.clone()
    sfa2[:valid] -= 8  # exponent - 1: every A scale halves exactly
    y2 = agemm.matmul(qa, qb, sfa2, sfb, 2e-4)
    torch.cuda.synchronize()
    assert torch.equal(y1, y2)


def test_tensor_scale_is_read_on_the_device_and_captures_in_a_cuda_graph():
    """`scale` given as a 1-element CUDA tensor must reach the kernel as a device pointer: no .item() (host sync), so the
    call can be captured in a CUDA graph -- and the replay must see a NEW value written to the tensor afterwards.  (The
    reference converts the tensor to `const float` on the host, kernels/src/bindings.cpp:104.)"""
    from arcquant_b200 import agemm
    M, N, K = 256, 512, 1088
    g = torch.Generator(device="cuda").manual_seed(3)
    qa = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, device="cuda", generator=g)
    qb = torch.randint(0, 256, (N, K // 2), dtype=torch.uint8, device="cuda", generator=g)
    sfa = torch.randint(0x30, 0x40, (agemm.sf_bytes(M, K),), dtype=torch.uint8, device="cuda", generator=g)
    sfb = torch.randint(0x30, 0x40, (agemm.sf_bytes(N, K),), dtype=torch.uint8, device="cuda", generator=g)
    scale = torch.tensor([1e-3], dtype=torch.float32, device="cuda")
    out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    agemm.matmul(qa, qb, sfa, sfb, scale, out=out)          # warm-up outside the capture (tensor maps, attributes)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            agemm.matmul(qa, qb, sfa, sfb, scale, out=out)  # a host read of `scale` here would invalidate the capture
    torch.cuda.current_stream().wait_stream(side)
    scale.fill_(2e-3)
    graph.replay()
    torch.cuda.synchronize()
    want = agemm.matmul(qa, qb, sfa, sfb, 2e-3)
    torch.cuda.synchronize()
    assert torch.equal(out, want)
