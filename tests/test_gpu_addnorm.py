"""arcq_add_rmsnorm_quantize_x: the residual add of the decoder layer fused in front of the norm +
quantise of the next block -- `hidden_states = residual + hidden_states` then `rmsnorm_quantize_x`
(/root/reference/benchmarks/modeling_arc.py:294-301, model/qLlamaLayer.py:162-166,295-300).  Bit-exact contract:
  sum_out == torch's bf16 `residual + x`            (fp32 add, round to nearest even)
  codes / scales == rmsnorm_quantize_x(sum_out)     (itself pinned to the reference's kernel by test_gpu_refkernels.py and to
                                                     the CPU oracle by test_gpu_quant.py)
for the shape-driven dispatch, every forced tile organisation, ragged M, and values that cancel / overflow in the add.
"""
import numpy as np
import pytest
import torch

import oracle
from tests._util import bits

pytestmark = pytest.mark.gpu


def _inputs(M, H, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(M, H, generator=g) * 3).to(torch.bfloat16)
    r = (torch.randn(M, H, generator=g) * 5).to(torch.bfloat16)
    r[::7, ::5] = -x[::7, ::5]                       # exact cancellation (+0)
    r[1::11, 3::9] = (x[1::11, 3::9].float() * -1.0078125).to(torch.bfloat16)   # near cancellation: the sum needs all fp32 bits
    x[2::13, 1::17] *= 300.0                         # large against small: the add rounds
    w = (1.0 + 0.1 * torch.randn(H, generator=g)).to(torch.bfloat16)
    idx = torch.randperm(H, generator=g).to(torch.int16)
    return x.cuda(), r.cuda(), w.cuda(), idx.cuda()


CASES = [("", 8192, 4096, 256), ("", 1300, 3584, 256), ("", 2500, 8192, 512), ("", 700, 2048, 64), ("", 129, 5120, 320), ("", 77, 4096, 0),
         ("8,512,1,8,1", 260, 2048, 128), ("8,512,1,4,1", 520, 2048, 128), ("4,256,1,4,2", 333, 2048, 128), ("4,512,1,4,1", 300, 4096, 256),
         ("2,512,1,2,1", 300, 8192, 512), ("2,256,2,2,2", 270, 2048, 64), ("1,512,2,1,1", 200, 4096, 256), ("4,512,2,4,1", 1000, 3584, 256)]


@pytest.mark.parametrize("knob,M,H,KE", CASES)
def test_fused_add_equals_torch_add_then_rmsnorm_quantize(arcq_option, knob, M, H, KE):
    from arcquant_b200 import agemm
    x, r, w, idx = _inputs(M, H, M + H)
    if knob:
        arcq_option("quant_tile", knob)
    q, sf, s = agemm.add_rmsnorm_quantize_x(x, r, w, 1e-6, idx, KE)
    s_ref = r + x                                      # torch's bf16 add
    q_ref, sf_ref = agemm.rmsnorm_quantize_x(s_ref, w, 1e-6, idx, KE)
    torch.cuda.synchronize()
    assert torch.equal(s.view(torch.int16), s_ref.view(torch.int16)), "sum differs from torch's bf16 add (%s)" % (knob or "default")
    off = torch.from_numpy(oracle.lib().sf_valid_offsets(M, H + KE)).cuda()
    assert torch.equal(q, q_ref), "codes differ from rmsnorm_quantize_x(residual + x) (%s)" % (knob or "default")
    assert torch.equal(sf[off], sf_ref[off]), "scales differ (%s)" % (knob or "default")


def test_fused_add_against_the_cpu_oracle():
    """Independent of our own rmsnorm kernel: numpy add (fp32, RNE to bf16) + the C restatement of rmsnorm.cu, allowing
    the handful of codes an rsqrtf difference of <= 2 ulp between CPU and GPU can move (as tests/test_gpu_quant.py does)."""
    from arcquant_b200 import agemm
    O = oracle.lib()
    M, H, KE = 96, 4096, 256
    x, r, w, idx = _inputs(M, H, 11)
    q, sf, s = agemm.add_rmsnorm_quantize_x(x, r, w, 1e-6, idx, KE)
    torch.cuda.synchronize()
    h = oracle.f32_to_bf16_bits(oracle.bf16_bits_to_f32(bits(x)) + oracle.bf16_bits_to_f32(bits(r)))
    assert np.array_equal(bits(s), h), "sum differs from the numpy restatement"
    q_o, sf_o, _ = O.rmsnorm_quantize_x(h, bits(w), 1e-6, idx.cpu().numpy(), KE)
    off = O.sf_valid_offsets(M, H + KE)
    nq = int((q.cpu().numpy() != q_o).sum())
    nsf = int((sf.cpu().numpy()[off] != sf_o[off]).sum())
    # rsqrtf on the CPU vs the GPU may differ in the last ulps: allow the handful of codes that sit on a rounding boundary
    assert nq <= q_o.size // 2000 and nsf <= max(2, sf_o[off].size // 2000), (nq, nsf)


def test_fused_add_rejects_bad_arguments():
    from arcquant_b200 import agemm
    x, r, w, idx = _inputs(64, 2048, 3)
    with pytest.raises(RuntimeError):
        agemm.add_rmsnorm_quantize_x(x, r[:32], w, 1e-6, idx, 64)
    with pytest.raises(RuntimeError):
        agemm.add_rmsnorm_quantize_x(x[:, :1024].contiguous(), r[:, :1024].contiguous(), w[:1024], 1e-6, idx[:1024], 64)   # H < 2048
