"""Host-side bookkeeping of the fused launches, on the CPU (no kernels): concatenating / interleaving already-quantised
weights must move whole 128-row scale-factor blocks, so that dequantising the fused operand gives back the parts."""
import numpy as np
import pytest
import torch

import oracle
from arcquant_b200 import agemm
from tests._util import bits, make_acts, make_weights


def _quant_w(O, N, KQ, KE, idx, seed):
    w, sw = make_weights(N, KQ, seed=seed)
    q, sf = O.quantize_w(bits(w), idx.numpy(), KE)
    return torch.from_numpy(q), torch.from_numpy(sf), sw


def test_fuse_weights_moves_whole_scale_blocks():
    O = oracle.lib()
    KQ, KE = 512, 64
    K = KQ + KE
    _, idx = make_acts(8, KQ, KE, seed=1)
    parts = [_quant_w(O, n, KQ, KE, idx, 10 + i) for i, n in enumerate((256, 128, 384))]
    qw, sfw, cs = agemm.fuse_weights(parts)
    assert qw.shape == (768, K // 2) and sfw.numel() == (768 // 128 + 1) * 128 * K // 16 and cs.shape == (768,)
    deq = O.dequant(qw.numpy(), sfw.numpy(), K)
    n0 = 0
    for q, sf, sw in parts:
        n = q.shape[0]
        assert np.array_equal(deq[n0:n0 + n], O.dequant(q.numpy(), sf.numpy(), K))
        assert torch.all(cs[n0:n0 + n] == torch.tensor(sw, dtype=torch.float32))
        n0 += n
    with pytest.raises(RuntimeError):
        agemm.fuse_weights([_quant_w(O, 200, KQ, KE, idx, 3), parts[0]])   # 200 rows: not whole scale blocks


def test_interleave_gate_up_layout():
    O = oracle.lib()
    KQ, KE, I = 256, 64, 384
    K = KQ + KE
    _, idx = make_acts(8, KQ, KE, seed=2)
    gate, up = _quant_w(O, I, KQ, KE, idx, 20), _quant_w(O, I, KQ, KE, idx, 21)
    qw, sfw, cs = agemm.interleave_gate_up(gate, up)
    assert qw.shape == (2 * I, K // 2) and cs.shape == (2 * I,)
    deq = O.dequant(qw.numpy(), sfw.numpy(), K)
    dg, du = O.dequant(gate[0].numpy(), gate[1].numpy(), K), O.dequant(up[0].numpy(), up[1].numpy(), K)
    for j in range(I // 128):   # rows [256 j, +128) = gate block j, rows [256 j + 128, +128) = up block j
        assert np.array_equal(deq[256 * j:256 * j + 128], dg[128 * j:128 * j + 128])
        assert np.array_equal(deq[256 * j + 128:256 * j + 256], du[128 * j:128 * j + 128])
        assert torch.all(cs[256 * j:256 * j + 128] == torch.tensor(gate[2], dtype=torch.float32))
        assert torch.all(cs[256 * j + 128:256 * j + 256] == torch.tensor(up[2], dtype=torch.float32))


def test_bench_layer_items_and_sampling():
    import bench
    items = bench.layer_items("llama3-8b")
    assert [(n, kq, ke) for n, kq, ke, _ in items] == [("qkv", 4096, 256), ("o", 4096, 256), ("gate_up", 4096, 256), ("down", 14336, 512)]
    assert sum(2.0 * 8192 * n * (kq + ke) for _, kq, ke, ns in items for n in ns) == pytest.approx(3.771e12, rel=1e-3)
    assert bench.quant_bytes(8192, 4096, 256) == 8192 * 10640 + 2 * 4096
    assert bench.layer_items("llama3-70b")[3][1] == 28672


# ---- placement plans (csrc/plan.cu): host-only index arithmetic, so the whole thing is checkable without a GPU ----------

def _plan(idx):
    import ctypes
    from arcquant_b200 import _lib
    L = _lib.load()
    KQ = idx.size
    plan = np.zeros(int(L.arcq_plan_entries(KQ)), dtype=np.uint16)
    rc = L.arcq_plan_build(idx.ctypes.data_as(ctypes.c_void_p), KQ, plan.ctypes.data_as(ctypes.c_void_p))
    return L, rc, plan


def _conflict(L, idx, R, plan):
    import ctypes
    return int(L.arcq_plan_max_conflict(idx.ctypes.data_as(ctypes.c_void_p), idx.size, R, plan.ctypes.data_as(ctypes.c_void_p)))


@pytest.mark.parametrize("KQ", [128, 512, 2048, 3584, 4096, 5120, 8192, 11008, 14336, 18944, 28672])
@pytest.mark.parametrize("kind", ["random", "identity", "reversed", "argsort"])
def test_plan_is_a_conflict_free_bijection(KQ, kind):
    rng = np.random.default_rng(KQ)
    idx = {"random": rng.permutation(KQ), "identity": np.arange(KQ), "reversed": np.arange(KQ)[::-1],
           "argsort": np.argsort(rng.standard_normal(KQ) * (1 + 20 * (rng.random(KQ) < 0.06)), kind="stable")}[kind].astype(np.int16)
    idx = np.ascontiguousarray(idx)
    L, rc, plan = _plan(idx)
    assert rc == 0, L.arcq_last_error()
    for s, R in enumerate((8, 4, 2)):
        sec = plan[2 * s * KQ:(2 * s + 1) * KQ]
        assert np.array_equal(np.sort(sec), np.arange(KQ)), "section R=%d is not a bijection" % R
        usable = KQ % (16 * 64 // R) == 0
        assert _conflict(L, idx, R, plan) == (1 if usable else -1)
    # what the plan removes: the identity placement of a random permutation is far from conflict-free
    if kind == "random" and KQ % 512 == 0:
        ident = np.tile(np.arange(KQ, dtype=np.uint16), 6)
        assert _conflict(L, idx, 2, ident) >= 3


def test_plan_build_rejects_a_non_permutation():
    idx = np.arange(256, dtype=np.int16)
    idx[5] = 6
    L, rc, _ = _plan(np.ascontiguousarray(idx))
    assert rc != 0 and b"permutation" in L.arcq_last_error()


# ---- calibration artefacts in the reference's on-disk schema (arcquant_b200/calib.py) --------------------------------

def test_calibration_artifacts_roundtrip_in_the_reference_schema(tmp_path):
    from arcquant_b200 import calib as C
    ri, sn, sc = C.synthetic_artifacts(hidden=512, intermediate=1536, n_layers=2, seed=3)
    # the reference's own lookups: model/qLlamaLayer.py:237-261, benchmarks/modeling_arc.py:95-106, model_utils.py:38-42
    nameTemplate = 'layers.{}.{}.{}.{}'
    for i in range(2):
        for block, projs in (("self_attn", C.ATTN_PROJS), ("mlp", C.MLP_PROJS)):
            for proj in projs:
                k = nameTemplate.format(i, block, proj, 'input')
                assert k in ri and k in sn and sn[k] % 64 == 0 and ri[k].dtype == torch.int64
        assert torch.equal(ri[C.key(i, "self_attn", "q_proj")], ri[C.key(i, "self_attn", "k_proj")])   # shared input
        assert ri[C.key(i, "mlp", "down_proj")].numel() == 1536
    C.validate(ri, sn)
    # outliers last (utilize.py:42): the last select_num reordered channels carry the largest statistics
    k = C.key(0, "mlp", "down_proj")
    assert sc[k][ri[k][-sn[k]:]].min() >= sc[k][ri[k][:-sn[k]]].max()
    paths = C.save_artifacts("Llama-Tiny", ri, sn, sc, root=str(tmp_path))
    assert paths["reorder_index"].endswith("llama-tiny_reorder_index_wikitext2_max.pt")   # reorder_indices.py:88
    ri2, sn2, sc2 = C.load_artifacts("llama-tiny", root=str(tmp_path))
    assert sn2 == sn and all(torch.equal(ri2[k], ri[k]) for k in ri) and all(torch.equal(sc2[k], sc[k]) for k in sc)
    e = C.layer_entries(ri2, sn2, 1)
    assert e["qkv"][0].dtype == torch.int16 and e["down"][1] == sn[C.key(1, "mlp", "down_proj")]
    with pytest.raises(FileNotFoundError):
        C.load_artifacts("absent-model", root=str(tmp_path))
    bad = dict(ri)
    bad[k] = ri[k].clone()
    bad[k][0] = bad[k][1]
    with pytest.raises(ValueError):
        C.validate(bad, sn)
    with pytest.raises(ValueError):
        C.validate(ri, {**sn, k: 100})
    assert C.select_num_for(4096, 0.06) == 256 and C.select_num_for(4096, 0.0) == 0


def test_peer_constants_match_the_header_and_bad_arguments_are_rejected_without_a_gpu():
    """arcquant_b200/peer.py mirrors the limits of include/arcq.h; the argument checks of the peer entry points run before
    anything touches a device (no launch happens for a rejected call)."""
    import ctypes
    import os
    import re
    from arcquant_b200 import _lib, peer as P
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "arcq.h")).read()
    macro = lambda n: int(re.search(r"#define\s+%s\s+(\d+)" % n, hdr).group(1))
    assert (P.MAX_PEERS, P.FLAG_BYTES, P.CHANNELS) == (macro("ARCQ_MAX_PEERS"), macro("ARCQ_PEER_FLAG_BYTES"), macro("ARCQ_PEER_CHANNELS"))
    assert macro("ARCQ_PEER_HANDLE_BYTES") == 64
    assert P.world_bytes(8, 1024, 4096) == 8 * 1024 * 4096 * 2
    L = _lib.load()
    fake = (ctypes.c_void_p * 2)(0x10000, 0x20000)
    a = ctypes.c_void_p(0x1000)
    # rows_per_rank must be a multiple of 128 covering M
    assert L.arcq_gemm_nvfp4_scatter(a, a, a, a, 256, 128, 64, 1.0, None, None, fake, 2, 0, 64, 0, None) == -1
    assert "rows_per_rank" in _lib.last_error()
    assert L.arcq_gemm_nvfp4_scatter(a, a, a, a, 512, 128, 64, 1.0, None, None, fake, 2, 0, 128, 0, None) == -1
    # rank outside the world, too many peers, misaligned peer pointer
    assert L.arcq_peer_reduce(a, fake, 2, 2, 0, 8, 8, 0, a, 0, None) == -1
    assert L.arcq_peer_reduce(a, fake, 9, 0, 0, 8, 8, 0, a, 0, None) == -1
    odd = (ctypes.c_void_p * 2)(0x10000, 0x20004)
    assert L.arcq_peer_reduce(a, odd, 2, 0, 0, 8, 8, 0, a, 0, None) == -1 and "aligned" in _lib.last_error()
    # cols not a multiple of 8, channel out of range, segment sizes not multiples of 16
    assert L.arcq_peer_reduce(a, fake, 2, 0, 0, 8, 12, 0, a, 0, None) == -1
    assert L.arcq_peer_reduce(a, fake, 2, 0, P.CHANNELS, 8, 8, 0, a, 0, None) == -1
    assert L.arcq_peer_allgather(a, 40, fake, 0, None, 0, None, 0, fake, 2, 0, 0, None) == -1
    assert L.arcq_peer_allgather(a, 32, fake, 8, None, 0, None, 0, fake, 2, 0, 0, None) == -1
    # bias: alignment and the SwiGLU exclusion are checked before any launch
    assert L.arcq_gemm_nvfp4_ex2(a, a, a, a, 128, 256, 64, 1.0, None, None, ctypes.c_void_p(0x1008), 0, a, 0, None) == -1
    assert L.arcq_gemm_nvfp4_ex2(a, a, a, a, 128, 256, 64, 1.0, None, None, ctypes.c_void_p(0x2000), 1, a, 0, None) == -1
    assert "SwiGLU" in _lib.last_error()
    # fused residual add: null / misaligned residual
    i16 = ctypes.c_void_p(0x3000)
    assert L.arcq_add_rmsnorm_quantize_x(a, None, a, 1e-6, 4, 4096, i16, 256, a, a, a, None) == -1
    assert L.arcq_add_rmsnorm_quantize_x(a, a, a, 1e-6, 4, 1024, i16, 64, a, a, a, None) == -3   # H outside the reduction tree's range
