"""Tensor-parallel host logic (arcquant_b200/tp.py) on CPU: index arithmetic, and a world_size-2 gloo run in which
each rank quantises ITS shard with the CPU oracle (the checker; the product path is CUDA-only), multiplies, and the
partials are all-reduced -- the same data flow RowParallelQLinear runs with NCCL on GPUs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from arcquant_b200 import tp
from tests._util import bits, make_acts, make_weights


def test_split_rows():
    assert tp.split_rows(4096, 4, 1) == (1024, 2048)
    with pytest.raises(ValueError):
        tp.split_rows(4096 + 64, 4, 0, 64)


@pytest.mark.parametrize("KQ,KE,world", [(4096, 256, 2), (4096, 256, 8), (8192, 512, 4), (14336, 512, 8), (5120, 320, 8)])
def test_shard_reorder_index_properties(KQ, KE, world):
    g = torch.Generator().manual_seed(KQ + world)
    idx = torch.randperm(KQ, generator=g).to(torch.int16)
    top = set(idx[KQ - KE:].tolist())
    total = 0
    for r in range(world):
        local, ke = tp.shard_reorder_index(idx, KE, world, r)
        lo, hi = tp.split_rows(KQ, world, r, 64)
        assert local.dtype == torch.int16 and local.numel() == hi - lo
        assert sorted(local.tolist()) == list(range(hi - lo))                      # a permutation of the shard
        glob = [c for c in idx.tolist() if lo <= c < hi]
        assert [c + lo for c in local.tolist()] == glob                            # global order, restricted (stable)
        owned = len([c for c in glob if c in top])
        assert ke % 64 == 0 and owned <= ke <= hi - lo and ke - owned < 64         # rounded UP to 64
        assert set(c + lo for c in local.tolist()[hi - lo - owned:]) == top & set(range(lo, hi))  # outliers stay last
        total += ke
    assert total >= KE
    assert tp.shard_select_nums(idx, KE, world) == [tp.shard_reorder_index(idx, KE, world, r)[1] for r in range(world)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _shard_partial(O, x, w, idx, KE, world, rank):
    """One rank's partial product through the oracle: local permutation, local KE, own scales."""
    lo, hi = tp.split_rows(x.shape[1], world, rank, 64)
    lidx, ke = tp.shard_reorder_index(idx, KE, world, rank)
    xs, ws = x[:, lo:hi].contiguous(), w[:, lo:hi].contiguous()
    sx = float(tp.nvfp4_scale(xs, signed=False))
    sw = float(tp.nvfp4_scale(ws, signed=True))
    qa, sfa = O.quantize_x(bits((xs / sx).to(torch.bfloat16)), lidx.numpy(), ke, 0)
    qb, sfb = O.quantize_w(bits((ws / sw).to(torch.bfloat16)), lidx.numpy(), ke, 0)
    return O.gemm_fast(qa, qb, sfa, sfb, sx * sw)


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    O = oracle.lib()
    M, N, KQ, KE = 24, 40, 1024, 128
    x, idx = make_acts(M, KQ, KE, seed=5)
    w, _ = make_weights(N, KQ, seed=6)
    part = torch.from_numpy(_shard_partial(O, x, w, idx, KE, world, rank))
    y = part.clone()
    dist.all_reduce(y, op=dist.ReduceOp.SUM)
    # column-parallel: each rank owns N/world output columns of the same quantised activation; gather to check
    lo, hi = tp.split_rows(N, world, rank)
    sx = float(tp.nvfp4_scale(x, signed=False))
    sw = float(tp.nvfp4_scale(w[lo:hi], signed=True))
    qa, sfa = O.quantize_x(bits((x / sx).to(torch.bfloat16)), idx.numpy(), KE, 0)
    qb, sfb = O.quantize_w(bits((w[lo:hi] / sw).to(torch.bfloat16)), idx.numpy(), KE, 0)
    col = torch.from_numpy(O.gemm_fast(qa, qb, sfa, sfb, sx * sw))
    cols = [torch.empty_like(col) for _ in range(world)]
    dist.all_gather(cols, col)
    if rank == 0:
        np.savez(os.path.join(out_dir, "tp.npz"), y=y.numpy(), part0=part.numpy(), ycol=torch.cat(cols, 1).numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_row_and_column_parallel_world2_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    d = np.load(tmp_path / "tp.npz")
    O = oracle.lib()
    M, N, KQ, KE = 24, 40, 1024, 128
    x, idx = make_acts(M, KQ, KE, seed=5)
    w, _ = make_weights(N, KQ, seed=6)
    # (1) the all-reduced result is exactly the sum of the per-shard oracle partials
    want = sum(_shard_partial(O, x, w, idx, KE, world, r) for r in range(world))
    assert np.array_equal(d["y"], want)
    assert np.array_equal(d["part0"], _shard_partial(O, x, w, idx, KE, world, 0))
    # (2) and matches the unquantised fp32 product to NVFP4 quantisation noise (different 16-groups than 1 GPU)
    ref = x.float().numpy().astype(np.float64) @ w.float().numpy().astype(np.float64).T
    for got in (d["y"], d["ycol"]):
        rel = np.linalg.norm(got - ref) / np.linalg.norm(ref)
        assert rel < 0.12, rel


def test_row_slices_quantise_independently():
    """Sequence-parallel edge of bench.py's TP mode: every rank quantises its own M/world tokens and the codes + the
    128-row scale blocks are all-gathered.  That is exact because rows are independent and scale blocks are 128-row
    aligned: concatenating per-slice outputs reproduces quantising the whole matrix byte for byte."""
    O = oracle.lib()
    M, KQ, KE, world = 512, 1024, 128, 4
    x, idx = make_acts(M, KQ, KE, seed=9)
    K = KQ + KE
    q_full, sf_full = O.quantize_x(bits(x), idx.numpy(), KE, 0)
    qs, sfs = [], []
    for r in range(world):
        lo, hi = tp.split_rows(M, world, r, 128)
        q, sf = O.quantize_x(bits(x[lo:hi]), idx.numpy(), KE, 0)
        qs.append(q)
        sfs.append(sf[: (hi - lo) * K // 16])
    assert np.array_equal(np.concatenate(qs), q_full)
    assert np.array_equal(np.concatenate(sfs), sf_full[: M * K // 16])
