"""arcq_linear_noalloc (include/arcq.h; no counterpart in the reference):
quantise + GEMM of one linear with the intermediate NVFP4 activation in the library's own `static __device__` scratch
array, for callers that cannot allocate.  Checked through the C ABI against the existing entry points it is built from
(arcq_quantize_x into caller buffers, then arcq_gemm_nvfp4): the outputs must be byte-identical -- same kernels, same
dispatch -- on calls in a row with different shapes (the array keeps the previous call's contents), and the call must
leave the device's free memory untouched.  Collected last on purpose.
"""
import ctypes

import pytest
import torch

from tests._util import make_acts, make_weights

SCRATCH_BYTES = 16 * 1024 * 1024  # ARCQ_NOALLOC_SCRATCH_BYTES of include/arcq.h


def _lib():
    from arcquant_b200 import _lib as L
    return L, L.load()


def test_scratch_size_and_rejection_need_no_gpu():
    """The size rule is host arithmetic and the over-size rejection happens before anything touches a device."""
    L, lib = _lib()
    for M, KQ, KE in ((1, 4096, 256), (300, 4096, 256), (2003, 14336, 512), (6846, 4096, 256)):
        K = KQ + KE
        want = (M * (K // 2) + 255) // 256 * 256 + (M // 128 + 1) * 128 * K // 16
        assert lib.arcq_linear_noalloc_bytes(M, KQ, KE) == want <= SCRATCH_BYTES
    assert lib.arcq_linear_noalloc_bytes(6847, 4096, 256) > SCRATCH_BYTES
    assert lib.arcq_linear_noalloc_bytes(-1, 4096, 256) == -1
    rc = lib.arcq_linear_noalloc(None, 8192, 4096, None, 256, -1, None, None, 4096, 1.0, None, None, None)
    assert rc == -1  # ARCQ_EINVAL
    msg = lib.arcq_last_error().decode()
    assert "ARCQ_NOALLOC_SCRATCH_BYTES" in msg and str(SCRATCH_BYTES) in msg, msg


def _case(M, KQ, KE, N, seed):
    """(x, idx, qw, sfw, sw) on the GPU; the weight is quantised by the existing entry point."""
    from arcquant_b200 import agemm
    x, idx = make_acts(M, KQ, KE, seed=seed)
    w, sw = make_weights(N, KQ, seed=seed + 1)
    x, idx, w = x.cuda(), idx.cuda(), w.cuda()
    qw, sfw = agemm.reorder_quantize_w(w, idx, KE)
    return x, idx, qw, sfw, sw


def _existing(lib, L, x, idx, KE, qw, sfw, alpha):
    """arcq_quantize_x into caller-owned buffers, then arcq_gemm_nvfp4 -- what the new entry point must reproduce."""
    M, KQ = x.shape
    K, N = KQ + KE, qw.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    qx = torch.empty((M, K // 2), dtype=torch.uint8, device="cuda")
    sfx = torch.zeros((int(lib.arcq_sf_bytes(M, K)),), dtype=torch.uint8, device="cuda")
    y = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    L.check(lib.arcq_quantize_x(x.data_ptr(), M, KQ, idx.data_ptr(), KE, -1, qx.data_ptr(), sfx.data_ptr(), st), "arcq_quantize_x")
    L.check(lib.arcq_gemm_nvfp4(qx.data_ptr(), qw.data_ptr(), sfx.data_ptr(), sfw.data_ptr(), M, N, K, alpha, None, y.data_ptr(), st),
            "arcq_gemm_nvfp4")
    return y


def _noalloc(lib, L, x, idx, KE, qw, sfw, alpha, y):
    M, KQ = x.shape
    st = torch.cuda.current_stream().cuda_stream
    L.check(lib.arcq_linear_noalloc(x.data_ptr(), M, KQ, idx.data_ptr(), KE, -1, qw.data_ptr(), sfw.data_ptr(), qw.shape[0],
                                    alpha, None, y.data_ptr(), st), "arcq_linear_noalloc")


@pytest.mark.gpu
def test_linear_noalloc_equals_existing_entry_points_on_calls_in_a_row():
    L, lib = _lib()
    # big, then small (stale codes / scales of the first call stay behind the second one's), then a third shape whose
    # 128-row tail block is mostly padding, then the first again
    shapes = [(1500, 4096, 256, 1024, 11), (77, 2048, 128, 512, 12), (130, 14336, 512, 4096, 13), (1500, 4096, 256, 1024, 14)]
    for M, KQ, KE, N, seed in shapes:
        assert lib.arcq_linear_noalloc_bytes(M, KQ, KE) <= SCRATCH_BYTES
        x, idx, qw, sfw, sw = _case(M, KQ, KE, N, seed)
        alpha = float(sw) * 0.37
        want = _existing(lib, L, x, idx, KE, qw, sfw, alpha)
        y = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device="cuda")
        launches = lib.arcq_launch_count()
        _noalloc(lib, L, x, idx, KE, qw, sfw, alpha, y)
        assert lib.arcq_launch_count() - launches == 2, "one quantise + one GEMM launch of this library"
        torch.cuda.synchronize()
        assert torch.equal(y, want), "M=%d KQ=%d KE=%d N=%d: %d of %d outputs differ from arcq_quantize_x + arcq_gemm_nvfp4" % (
            M, KQ, KE, N, int((y != want).sum()), y.numel())
        assert bool(torch.isfinite(y.float()).all())
    # the same call twice in a row gives the same bytes, and allocates nothing: free device memory is unchanged
    y2 = torch.empty_like(y)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    _noalloc(lib, L, x, idx, KE, qw, sfw, alpha, y2)
    _noalloc(lib, L, x, idx, KE, qw, sfw, alpha, y2)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.equal(y2, want)
    assert free1 == free0 and torch.cuda.memory_reserved() == reserved0, "arcq_linear_noalloc changed the device's free memory"


@pytest.mark.gpu
def test_linear_noalloc_rejects_oversize_and_reads_alpha_from_the_device():
    L, lib = _lib()
    M, KQ, KE, N = 256, 4096, 256, 512
    x, idx, qw, sfw, sw = _case(M, KQ, KE, N, 21)
    st = torch.cuda.current_stream().cuda_stream
    y = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    big = torch.empty((8192, KQ), dtype=torch.bfloat16, device="cuda")
    rc = lib.arcq_linear_noalloc(big.data_ptr(), 8192, KQ, idx.data_ptr(), KE, -1, qw.data_ptr(), sfw.data_ptr(), N, 1.0, None,
                                 y.data_ptr(), st)
    assert rc == -1 and "ARCQ_NOALLOC_SCRATCH_BYTES" in lib.arcq_last_error().decode()
    # alpha_dev like arcq_gemm_nvfp4
    a = torch.tensor([float(sw) * 0.5], dtype=torch.float32, device="cuda")
    want = _existing(lib, L, x, idx, KE, qw, sfw, float(a))
    L.check(lib.arcq_linear_noalloc(x.data_ptr(), M, KQ, idx.data_ptr(), KE, -1, qw.data_ptr(), sfw.data_ptr(), N, 0.0,
                                    ctypes.c_void_p(a.data_ptr()), y.data_ptr(), st), "arcq_linear_noalloc")
    torch.cuda.synchronize()
    assert torch.equal(y, want)
