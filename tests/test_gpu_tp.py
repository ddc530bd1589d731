"""Row-/column-parallel NVFP4+ARC layers on one GPU: every rank's shard is run in turn (no collective needed to check
parity) and compared PER SHARD with the oracle on the same shard inputs."""
import numpy as np
import pytest
import torch

import oracle
from arcquant_b200 import tp
from tests._util import bits, make_acts, make_weights

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,KQ,KE,world", [(256, 512, 4096, 256, 2), (130, 256, 8192, 512, 8), (64, 384, 14336, 512, 4)])
def test_row_parallel_shards_match_oracle(M, N, KQ, KE, world):
    O = oracle.lib()
    x, idx = make_acts(M, KQ, KE, seed=KQ)
    w, _ = make_weights(N, KQ, seed=KQ + 1)
    total = np.zeros((M, N))
    for rank in range(world):
        layer = tp.RowParallelQLinear(w.cuda(), idx, KE, world, rank)
        lo, hi = layer.k_slice
        xs = x[:, lo:hi].contiguous().cuda()
        qx, sfx, sx = layer.quantize_input(xs)
        y = layer.forward_quantized(qx, sfx, sx)
        torch.cuda.synchronize()
        lidx, ke = tp.shard_reorder_index(idx, KE, world, rank)
        assert layer.select_num == ke and torch.equal(layer.reorder_index.cpu(), lidx)
        q_ref, sf_ref = O.quantize_x(bits((xs / sx).to(torch.bfloat16)), lidx.numpy(), ke)  # flavour by the KQ table, as agemm
        off = O.sf_valid_offsets(M, hi - lo + ke)
        assert np.array_equal(qx.cpu().numpy(), q_ref) and np.array_equal(sfx.cpu().numpy()[off], sf_ref[off])
        ref = O.gemm_fast(qx.cpu().numpy(), layer.W.cpu().numpy(), sfx.cpu().numpy(), layer.scale_w.cpu().numpy(),
                          float(sx) * float(layer.scale))
        got = y.float().cpu().numpy().astype(np.float64)
        assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 2e-3
        total += got
    full = x.float().numpy().astype(np.float64) @ w.float().numpy().astype(np.float64).T
    assert np.linalg.norm(total - full) / np.linalg.norm(full) < 0.12


def test_column_parallel_concat_equals_unsharded():
    """N-split: concatenated shard outputs vs the unsharded layer and the fp32 product."""
    M, N, KQ, KE, world = 128, 512, 4096, 256, 4
    x, idx = make_acts(M, KQ, KE, seed=1)
    w, _ = make_weights(N, KQ, seed=2)
    full = tp.QLinearShard(w.cuda(), idx, KE)
    qx, sfx, sx = full.quantize_input(x.cuda())
    y_full = full.forward_quantized(qx, sfx, sx)
    outs = []
    for rank in range(world):
        sh = tp.ColumnParallelQLinear(w.cuda(), idx, KE, world, rank)
        outs.append(sh.forward_quantized(qx, sfx, sx).float())
    y = torch.cat(outs, 1)
    # per-shard weight scales differ from the full-tensor scale, so codes differ: both must sit within NVFP4
    # quantisation noise of the unquantised product, and within twice that of each other
    ref = x.float().cuda() @ w.float().cuda().T
    assert (y - ref).norm() / ref.norm() < 0.12 and (y_full.float() - ref).norm() / ref.norm() < 0.12
    assert (y - y_full.float()).norm() / ref.norm() < 0.2
