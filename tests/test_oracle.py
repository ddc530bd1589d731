"""CPU tests that PIN the oracle (oracle/arcq_oracle.c, oracle/fakequant.py) before anything trusts it.

The reference ships no tests or golden vectors for this path, so the pins are outputs of the
reference's own code generated in the authoring container by tests/golden/make_golden.py:
  * cutlass_converters.npz -- the reference's vendored CUTLASS v4.3.1 host NumericConverters (the ones
    kernels/src/reorder.cu:103-109 instantiates): e2m1 of every bf16 value, e2m1 rounding boundaries, UE4M3 encode /
    decode, bf16 RNE;
  * cute_sf_layout.npz -- CuTe Sm1xxBlockScaledConfig<16>::tile_atom_to_shape_SFA offsets (reorder.cu:343);
  * fakequant_reference.npz -- /root/reference/model/quantize.py imported as is (fake_reorder_quantize_x / _w).
Nothing here needs a GPU or /root/reference.
"""
import os

import numpy as np
import pytest
import torch

import oracle
from oracle import fakequant as fq
from tests._util import bits, make_acts

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def O():
    return oracle.lib()


@pytest.fixture(scope="module")
def conv():
    return np.load(os.path.join(GOLD, "cutlass_converters.npz"))


def _f32(bits_u32):
    return np.asarray(bits_u32, dtype=np.uint32).view(np.float32)


# ---- scalar converters against the reference's CUTLASS host converters ---------------------------------

def test_e2m1_of_every_bf16(O, conv):
    tab = conv["e2m1_of_bf16"]
    bad = 0
    for h in range(65536):
        if tab[h] == 0xFF:  # NaN inputs were skipped by the probe
            continue
        x = float(_f32(np.uint32(h << 16)))
        t = min(6.0, max(-6.0, x))
        bad += int(O.L.arcq_o_f32_to_e2m1(t) != tab[h])
    assert bad == 0


def test_e2m1_boundaries_and_decode(O, conv):
    for u, want in zip(conv["e2m1_boundary_in"], conv["e2m1_boundary_out"]):
        assert O.L.arcq_o_f32_to_e2m1(float(_f32(u))) == want, hex(int(u))
    for c in range(16):
        got = O.L.arcq_o_e2m1_to_f32(c)
        assert np.float32(got).tobytes() == np.float32(conv["e2m1_decode"][c]).tobytes()  # -0.0 preserved


def test_ue4m3_encode_decode(O, conv):
    xs = _f32(conv["ue4m3_in"])
    want = conv["ue4m3_out"]
    got = np.array([O.L.arcq_o_f32_to_ue4m3(float(x)) for x in xs], dtype=np.uint8)
    assert (got == want).all(), "%d of %d UE4M3 encodes differ" % (int((got != want).sum()), len(xs))
    dec = np.array([O.L.arcq_o_ue4m3_to_f32(b) for b in range(127)], dtype=np.float32)
    assert (dec == conv["ue4m3_decode"]).all()
    # torch's float8_e4m3fn cast is the same RNE-saturating conversion on [2^-9, 448]
    t = torch.from_numpy(xs.copy()).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    assert (t == want).all()


def test_bf16_rne(O, conv):
    xs = _f32(conv["bf16_in"])
    got = np.array([O.L.arcq_o_f32_to_bf16(float(x)) for x in xs], dtype=np.uint16)
    assert (got == conv["bf16_out"]).all()
    assert (oracle.f32_to_bf16_bits(xs) == conv["bf16_out"]).all()


# ---- scale-factor layout against CuTe ------------------------------------------------------------------

@pytest.mark.parametrize("rows,K", [(300, 4160), (128, 64), (257, 1088)])
def test_sf_layout_matches_cute(O, rows, K):
    d = np.load(os.path.join(GOLD, "cute_sf_layout.npz"))
    want = d["off_%dx%d" % (rows, K)].astype(np.int64)
    got = O.sf_valid_offsets(rows, K)
    assert (got == want).all()
    for r, g in ((0, 0), (rows - 1, K // 16 - 1), (rows // 2, 3)):
        assert O.sf_offset(r, g, K) == want[r, g]
    # the reference's allocation (bindings.cpp:83-95) covers CuTe's cosize
    assert O.sf_bytes(rows, K) >= int(d["cosize_%dx%d" % (rows, K)][0])
    assert O.sf_bytes(rows, K) == (rows // 128 + 1) * 128 * K // 16
    assert len(np.unique(got)) == got.size  # a bijection onto distinct bytes


def test_flavour_table(O):
    # kernels/src/bindings.cpp:141-160
    for kq in (2048, 3072, 4096, 5120, 8192, 11008, 13824, 14336):
        assert O.flavour_for_kq(kq) == 0
    for kq in (3584, 18944, 27648, 28672):
        assert O.flavour_for_kq(kq) == 1
    assert O.flavour_for_kq(1024) == -1


# ---- the fake-quant restatement against the real model/quantize.py -------------------------------------

@pytest.mark.parametrize("tag", ["a", "b", "c", "d", "e"])
def test_fakequant_restatement_matches_reference_module(tag):
    """a-c: identity order (outliers last); d, e: a real permutation, so the reference's `reorder_index[-select_num:]`
    gather (model/quantize.py:237-241,262-267) is pinned too."""
    d = np.load(os.path.join(GOLD, "fakequant_reference.npz"))
    x = torch.from_numpy(oracle.bf16_bits_to_f32(d[tag + "_x_bits"]))
    KE = int(d[tag + "_KE"])
    idx = torch.from_numpy(d[tag + "_idx"].astype(np.int64)) if tag + "_idx" in d else torch.arange(x.shape[1])
    if tag in ("d", "e"):
        assert not torch.equal(idx, torch.arange(x.shape[1]))
    qx, _, scale = fq.fake_reorder_quantize_x(x, idx, KE)
    assert np.float32(float(scale)) == d[tag + "_scale"]
    assert np.array_equal(qx.numpy().astype(np.float32), d[tag + "_qx"])
    w = torch.from_numpy(d[tag + "_w"])
    qw, _, sw = fq.fake_reorder_quantize_w(w, idx, KE)
    assert np.float32(float(sw)) == d[tag + "_sw"]
    assert np.array_equal(qw.numpy().astype(np.float32), d[tag + "_qw"])


# ---- kernel-faithful oracle vs fake-quant: scales exact, codes exact modulo the documented tie rule ----

def slot_maps(KQ, KE, flavour):
    """Slot (16-group position in the K = KQ+KE layout) of every primary group and of every residual group
    (reorder.cu:139,175 16-wide; reorder.cu:451-463,510-515 32-wide)."""
    G = (KQ - KE) // 16
    prim = np.arange(KQ // 16)
    res = np.zeros(KE // 16, dtype=np.int64)
    for j in range(KE // 16):
        if flavour == 0:
            p = G + 2 * j
            r = p + 1
        else:
            p = G + 4 * (j >> 1) + (j & 1)
            r = p + 2
        prim[G + j] = p
        res[j] = r
    return prim, res


def split_layout(a, KQ, KE, flavour):
    """[rows, KQ+KE] in kernel layout -> ([rows, KQ] primary in reordered order, [rows, KE] residual)."""
    prim, res = slot_maps(KQ, KE, flavour)
    g = a.reshape(a.shape[0], -1, 16)
    return g[:, prim].reshape(a.shape[0], KQ), g[:, res].reshape(a.shape[0], KE)


_TIES = (0.25, 0.75, 1.25, 1.75, 2.5, 3.5, 5.0)


@pytest.mark.parametrize("M,KQ,KE,flavour,dist", [(64, 1024, 128, 0, "A"), (32, 2048, 256, 0, "B"), (48, 512, 0, 0, "A"),
                                                  (32, 1024, 128, 1, "A")])
def test_oracle_vs_fakequant_modulo_ties(O, M, KQ, KE, flavour, dist):
    x, _ = make_acts(M, KQ, KE, seed=7 + KQ, dist=dist)
    order = torch.argsort(x.float().abs().amax(0), stable=True)
    x = x[:, order].contiguous()  # identity index, outliers last: fake-quant and the kernels form the same groups
    xf = x.float()
    r, c = np.unravel_index(int(torch.argmax(xf.abs())), xf.shape)
    x[r, c] = 2688.0  # signed max exactly 448*6 -> the fake path's own per-tensor scale is exactly 1
    xf = x.float()
    idx = torch.arange(KQ)
    fake, _, scale = fq.fake_reorder_quantize_x(xf, idx, KE)
    assert float(scale) == 1.0
    q, sf = O.quantize_x(bits(x), idx.numpy().astype(np.int16), KE, flavour)
    deq = O.dequant(q, sf, KQ + KE)
    prim, res = split_layout(deq, KQ, KE, flavour)
    sfv = np.array([O.L.arcq_o_ue4m3_to_f32(int(b)) for b in range(127)], dtype=np.float32)
    sf_rows = sfv[np.minimum(sf[O.sf_valid_offsets(M, KQ + KE)], 126)]  # [M, K/16] decoded scales
    sf_prim, _ = split_layout(np.repeat(sf_rows, 16, axis=1), KQ, KE, flavour)
    # fake-quant's scale for the same groups
    g = xf.view(M, KQ // 16, 16)
    fsf = fq.quantize_ue4m3((g.abs().amax(-1) / 6.0).clamp_min(1e-9)).numpy()
    live = g.abs().amax(-1).numpy() / 6.0 >= 2.0 ** -6  # above E4M3's normal range floor (fake has no subnormals)
    assert (np.repeat(fsf, 16, 1)[np.repeat(live, 16, 1)] == sf_prim[np.repeat(live, 16, 1)]).all(), "scales differ"
    fprim = fake[:, :KQ].numpy()
    diff = (fprim != prim) & np.repeat(live, 16, 1)
    ratio = np.abs(xf.numpy()) / sf_prim
    recip = np.abs(xf.numpy() * (np.float32(1.0) / sf_prim).astype(np.float32))
    near_tie = np.zeros_like(diff)
    for t in _TIES:
        near_tie |= (np.abs(ratio - t) <= 1e-6 * t) | (np.abs(recip - t) <= 1e-6 * t)
    assert not (diff & ~near_tie).any(), "%d code mismatches that are not rounding ties" % int((diff & ~near_tie).sum())
    assert diff.mean() < 0.02
    if KE and flavour == 0:  # (the 32-wide kernels subtract with the UNROUNDED scale, reorder.cu:474: no fake twin)
        # the residual is requantised from bf16(x - deq) in the kernels and from fp32 (x - deq) in fake-quant: same
        # values except where bf16 rounding moves an amax or a code across a boundary
        fres = fake[:, KQ:].numpy()
        assert (fres != res).mean() < 0.05


def test_oracle_w_duplicates_primary(O):
    """Weights: ARC slots hold a copy of the primary codes and scale (reorder.cu:307-316)."""
    KQ, KE = 1024, 128
    g = torch.Generator().manual_seed(0)
    w = (torch.randn(40, KQ, generator=g) * 30).to(torch.bfloat16)
    idx = torch.randperm(KQ, generator=g).to(torch.int16).numpy()
    for flavour in (0, 1):
        q, sf = O.quantize_w(bits(w), idx, KE, flavour)
        deq = O.dequant(q, sf, KQ + KE)
        prim, res = split_layout(deq, KQ, KE, flavour)
        assert np.array_equal(prim[:, KQ - KE:], res)


def test_oracle_x_residual_improves_error(O):
    """ARC property: primary + residual reconstructs the outlier channels better than the primary alone."""
    M, KQ, KE = 32, 1024, 128
    x, idx = make_acts(M, KQ, KE, seed=3)
    q, sf = O.quantize_x(bits(x), idx.numpy(), KE, 0)
    prim, res = split_layout(O.dequant(q, sf, KQ + KE), KQ, KE, 0)
    xr = x.float().numpy()[:, idx.numpy().astype(np.int64)]
    e1 = np.abs(xr[:, KQ - KE:] - prim[:, KQ - KE:]).mean()
    e2 = np.abs(xr[:, KQ - KE:] - prim[:, KQ - KE:] - res).mean()
    assert e2 < 0.25 * e1


def test_oracle_gemm_matches_numpy(O):
    M, N, KQ, KE = 9, 7, 256, 64
    x, idx = make_acts(M, KQ, KE, seed=1)
    w, _ = make_acts(N, KQ, KE, seed=2)
    qa, sfa = O.quantize_x(bits(x), idx.numpy(), KE, 0)
    qb, sfb = O.quantize_w(bits(w), idx.numpy(), KE, 0)
    ref = O.gemm(qa, qb, sfa, sfb, 0.5)
    fast = O.gemm_fast(qa, qb, sfa, sfb, 0.5)
    assert np.allclose(ref, fast, rtol=1e-12, atol=0)


def test_oracle_rejects_bad_shapes(O):
    x = np.zeros((2, 100), dtype=np.uint16)
    with pytest.raises(ValueError):
        O.quantize_x(x, np.arange(100, dtype=np.int16), 0, 0)


def test_div6_exact():
    """The kernels' division-free amax/6 (q0 = a*r6; rem = fma(-6, q0, a); q = fma(rem, r6, q0), quant_ws.cu div6) is
    bit-identical to the correctly rounded fp32 division (reorder.cu:138) for EVERY finite non-negative bf16 amax."""
    from fractions import Fraction

    def rn32(fr):
        if fr == 0:
            return np.float32(0)
        f = np.float32(float(fr))
        cands = [np.nextafter(f, np.float32(-np.inf)), f, np.nextafter(f, np.float32(np.inf))]
        return np.float32(min(cands, key=lambda c: (abs(Fraction(float(c)) - fr), int(np.float32(c).view(np.uint32)) & 1)))

    r6 = np.float32(1.0) / np.float32(6.0)
    assert int(r6.view(np.uint32)) == 0x3E2AAAAB
    R = Fraction(float(r6))
    for h in list(range(0, 0x7F80, 7)) + list(range(0x4000, 0x4600)):
        a = Fraction(float(np.array([h << 16], dtype=np.uint32).view(np.float32)[0]))
        q0 = Fraction(float(rn32(a * R)))
        rem = Fraction(float(rn32(a - 6 * q0)))
        q = rn32(q0 + rem * R)
        assert q.view(np.uint32) == rn32(a / 6).view(np.uint32), hex(h)


# ---- the torch sampled-row GEMM oracle (oracle/torchref.py) against the C oracle --------------------------------

def test_torchref_matches_c_oracle(O):
    from oracle import torchref as T
    rng = np.random.default_rng(5)
    M, N, K = 200, 136, 320
    qa = rng.integers(0, 256, (M, K // 2), dtype=np.uint8)
    qb = rng.integers(0, 256, (N, K // 2), dtype=np.uint8)
    sfa = rng.integers(0, 0x7f, O.sf_bytes(M, K), dtype=np.uint8)   # includes E4M3 subnormals (exponent field 0)
    sfb = rng.integers(0x20, 0x50, O.sf_bytes(N, K), dtype=np.uint8)
    rows = torch.tensor([0, 1, 31, 32, 127, 128, 129, 199])
    da = T.dequant_rows(torch.from_numpy(qa), torch.from_numpy(sfa), rows, K).numpy()
    assert np.array_equal(da, O.dequant(qa, sfa, K)[rows.numpy()].astype(np.float64))
    ref, mag = T.gemm_rows(torch.from_numpy(qa), torch.from_numpy(sfa), torch.from_numpy(qb), torch.from_numpy(sfb), rows, 0.37)
    full = O.gemm(qa, qb, sfa, sfb, 0.37)
    assert np.allclose(ref.numpy(), full[rows.numpy()], rtol=1e-12, atol=0)
    assert (mag.numpy() >= np.abs(ref.numpy()) - 1e-9).all()
    assert np.array_equal(T.sf_offsets(torch.arange(M), K).numpy(), O.sf_valid_offsets(M, K))
    r = T.sample_rows(1000)
    assert r[0] == 0 and r[-1] == 999 and all(((r >= b * 128) & (r < b * 128 + 128)).sum() >= 1 for b in range(8))


def test_bf16_add_restatement_is_torchs_bf16_add():
    """The fused residual add (arcq_add_rmsnorm_quantize_x) and the bias epilogue (arcq_gemm_nvfp4_ex2) are specified as
    torch's bf16 `a + b`: an fp32 add rounded to nearest even.  tests/test_gpu_addnorm.py checks the kernel against the
    numpy restatement `f32_to_bf16_bits(f32(a) + f32(b))`; this pins that restatement to torch itself on the CPU, on random
    values, exact cancellations, sums that need more than 8 bits, ties, overflow to inf and subnormals."""
    import torch
    g = torch.Generator().manual_seed(0)
    a = (torch.randn(1 << 16, generator=g) * 8).to(torch.bfloat16)
    b = (torch.randn(1 << 16, generator=g) * 8).to(torch.bfloat16)
    b[::5] = -a[::5]
    b[1::7] = (a[1::7].float() * -1.0078125).to(torch.bfloat16)
    a[2::11] *= 2.0 ** 20
    a[3::13] = torch.tensor(3.3895e38).to(torch.bfloat16)          # near the top of the range: a + a overflows
    b[3::13] = a[3::13]
    a[4::17] = torch.tensor(1e-39).to(torch.bfloat16)              # subnormal operands
    # every pair of the 2^8 x 2^8 mantissa/exponent neighbourhood around 1.0: all rounding ties of the add
    lo = torch.arange(0x3f00, 0x4000, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)
    a = torch.cat((a, lo.repeat_interleave(lo.numel())))
    b = torch.cat((b, (lo.float() * 2.0 ** -7).to(torch.bfloat16).repeat(lo.numel())))
    want = (a + b).view(torch.int16).numpy().view(np.uint16)
    ab = a.view(torch.int16).numpy().view(np.uint16)
    bb = b.view(torch.int16).numpy().view(np.uint16)
    with np.errstate(over="ignore"):
        got = oracle.f32_to_bf16_bits(oracle.bf16_bits_to_f32(ab) + oracle.bf16_bits_to_f32(bb))
    nan = np.isnan(oracle.bf16_bits_to_f32(want))
    assert np.array_equal(got[~nan], want[~nan])
    assert np.isnan(oracle.bf16_bits_to_f32(got[nan])).all()
