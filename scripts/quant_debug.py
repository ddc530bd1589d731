"""Classify mismatches of the quantise kernel against the CPU oracle (GPU box, debugging aid)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle
from arcquant_b200 import agemm
from tests._util import bits, make_acts
from tests.test_oracle import slot_maps

O = oracle.lib()
for (M, KQ, KE) in [(128, 4096, 256), (130, 14336, 512), (64, 3584, 256), (20, 28672, 1024), (129, 8192, 512), (48, 640, 64)]:
    x, idx = make_acts(M, KQ, KE, seed=M + KQ)
    fl = max(O.flavour_for_kq(KQ), 0)
    q_ref, sf_ref = O.quantize_x(bits(x), idx.numpy(), KE)
    q, sf = agemm.reorder_quantize_x(x.cuda(), idx.cuda(), KE)
    torch.cuda.synchronize()
    q, sf = q.cpu().numpy(), sf.cpu().numpy()
    K = KQ + KE
    off = O.sf_valid_offsets(M, K)
    prim, res = slot_maps(KQ, KE, fl)
    dq = (q != q_ref).reshape(M, K // 16, 8).any(-1)   # [M, slots]
    ds = sf[off] != sf_ref[off]
    G = (KQ - KE) // 16
    print("M=%d KQ=%d KE=%d flavour=%d: codes bad plain=%d arc-primary=%d residual=%d | sf bad plain=%d arc-primary=%d residual=%d | rows bad=%s" % (
        M, KQ, KE, fl, dq[:, prim[:G]].sum(), dq[:, prim[G:]].sum(), dq[:, res].sum() if KE else 0,
        ds[:, prim[:G]].sum(), ds[:, prim[G:]].sum(), ds[:, res].sum() if KE else 0,
        np.nonzero(dq.any(1) | ds.any(1))[0][:10]), flush=True)
