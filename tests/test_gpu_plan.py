"""Conflict-free placement plans (csrc/plan.cu, arcq_quantize_x_ex): the planned kernels must produce the oracle's bytes,
with the right plan, with plans disabled, and with a plan built for a DIFFERENT reorder_index (a plan is only ever an
optimisation: any bijection is a correct placement).  Reference semantics: kernels/src/reorder.cu:68-203,380-555."""
import numpy as np
import pytest
import torch

import oracle
from tests._util import bits, make_acts

pytestmark = pytest.mark.gpu


def _same(q, sf, q_ref, sf_ref, M, K):
    off = oracle.lib().sf_valid_offsets(M, K)
    return int((q.cpu().numpy() != q_ref).sum()) + int((sf.cpu().numpy()[off] != sf_ref[off]).sum())


CASES = [  # knob (R,threads,S,RT,CTAs/SM), M, KQ, KE
    ("8,512,1,4,1", 1290, 4096, 256), ("8,512,1,8,1", 260, 4096, 256), ("8,512,1,4,1", 200, 3584, 256), ("8,512,1,4,1", 300, 2048, 0),
    ("4,256,2,4,2", 1300, 4096, 256), ("4,512,2,4,1", 700, 8192, 512), ("4,256,2,4,2", 260, 3584, 256), ("4,512,1,4,1", 200, 5120, 320),
    ("2,512,1,2,1", 600, 14336, 512), ("2,512,1,2,1", 300, 18944, 512), ("2,256,2,2,2", 333, 4096, 256), ("2,256,2,2,2", 150, 3584, 256),
    ("4,512,1,4,1", 140, 11008, 1024),   # 11008 = 43 x 256: usable for R = 4, not for R = 2
]


@pytest.mark.parametrize("knob,M,KQ,KE", CASES)
def test_planned_kernels_match_oracle(arcq_option, knob, M, KQ, KE):
    from arcquant_b200 import agemm
    if agemm.BACKEND != "pybind11":
        pytest.skip("plans are built by the compiled binding")
    O = oracle.lib()
    x, idx = make_acts(M, KQ, KE, seed=M + KQ, dist="A")
    q_ref, sf_ref = O.quantize_x(bits(x), idx.numpy(), KE)
    arcq_option("quant_tile", knob)
    xd, idxd = x.cuda(), idx.cuda()
    R = int(knob.split(",")[0])
    plan = agemm.plan_build(idx)
    if KQ % (16 * 64 // R) == 0:
        assert agemm.plan_max_conflict(idx, R, plan) == 1
    try:
        agemm.use_plans(True)
        q1, sf1 = agemm.reorder_quantize_x(xd, idxd, KE)                  # plan from the binding's cache
        q2, sf2 = agemm.reorder_quantize_x_planned(xd, idxd, KE, plan.cuda())
        other = torch.randperm(KQ, generator=torch.Generator().manual_seed(1)).to(torch.int16)
        q3, sf3 = agemm.reorder_quantize_x_planned(xd, idxd, KE, agemm.plan_build(other).cuda())   # stale plan: still correct
        am = torch.tensor([1777.0], device="cuda")
        q5, sf5, _ = agemm.reorder_quantize_x_scaled(xd, idxd, KE, amax=am)   # planned + fused pre-scale
        agemm.use_plans(False)
        q4, sf4 = agemm.reorder_quantize_x(xd, idxd, KE)                  # arithmetic placement
        q6, sf6, _ = agemm.reorder_quantize_x_scaled(xd, idxd, KE, amax=am)   # (pinned to torch by tests/test_gpu_prescale.py)
    finally:
        agemm.use_plans(True)
    torch.cuda.synchronize()
    K = KQ + KE
    for tag, (q, sf) in {"cached plan": (q1, sf1), "explicit plan": (q2, sf2), "stale plan": (q3, sf3), "no plan": (q4, sf4)}.items():
        assert _same(q, sf, q_ref, sf_ref, M, K) == 0, "%s (%s M=%d KQ=%d KE=%d)" % (tag, knob, M, KQ, KE)
    off = torch.from_numpy(O.sf_valid_offsets(M, K)).cuda()
    assert torch.equal(q5, q6) and torch.equal(sf5[off], sf6[off]), "planned + pre-scale differs from unplanned + pre-scale"


@pytest.mark.parametrize("M,KQ,KE", [(8192, 4096, 256), (4096 + 77, 14336, 512), (2500, 8192, 512), (3000, 3584, 256)])
def test_default_dispatch_with_plans_at_prefill_sizes(M, KQ, KE):
    from arcquant_b200 import agemm
    O = oracle.lib()
    x, idx = make_acts(256, KQ, KE, seed=KQ, dist="A")
    reps = (M + 255) // 256
    x = x.repeat(reps, 1)[:M].contiguous()
    xd, idxd = x.cuda(), idx.cuda()
    q, sf = agemm.reorder_quantize_x(xd, idxd, KE)
    torch.cuda.synchronize()
    rows = torch.arange(0, M, max(1, M // 200))
    q_ref, sf_ref = O.quantize_x(bits(x[rows]), idx.numpy(), KE)
    from oracle import torchref as T
    K = KQ + KE
    assert np.array_equal(q[rows.cuda()].cpu().numpy(), q_ref)
    got_sf = sf[T.sf_offsets(rows.cuda(), K)].cpu().numpy()
    assert np.array_equal(got_sf, sf_ref[O.sf_valid_offsets(rows.numel(), K)])
