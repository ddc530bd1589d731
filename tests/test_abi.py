"""The C-ABI library loads and exports every symbol include/arcq.h declares (no compute calls: runs without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "arcq.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(arcq_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_hot_entry_points():
    names = _declared()
    for need in ("arcq_quantize_x", "arcq_quantize_w", "arcq_rmsnorm_quantize_x", "arcq_gemm_nvfp4", "arcq_sf_bytes",
                 "arcq_last_error"):
        assert need in names


def test_library_exports_every_declared_symbol():
    from arcquant_b200 import _lib
    L = _lib.load()
    assert os.path.dirname(_lib.lib_path()).startswith(ROOT)  # in-tree, not site-packages
    for name in _declared():
        assert hasattr(L, name), name
        assert name in _lib.SIGNATURES, "ctypes signature missing for " + name
    assert sorted(_lib.SIGNATURES) == _declared()


def test_host_only_entry_points():
    from arcquant_b200 import agemm
    # allocation size of the reference, kernels/src/bindings.cpp:83-95 (one extra 128-row tile, even when M%128==0)
    assert agemm.sf_bytes(2048, 4352) == (2048 // 128 + 1) * 128 * 4352 // 16
    assert agemm.sf_bytes(1, 64) == 128 * 4
    assert agemm.flavour_for_kq(4096) == 0 and agemm.flavour_for_kq(3584) == 1 and agemm.flavour_for_kq(1000) == -1
    L = __import__("arcquant_b200._lib", fromlist=["load"]).load()
    assert L.arcq_sf_offset(129, 5, 4160) == (4160 // 64) * 512 + 512 + 16 + 0 + 1
    assert b"arcq" in L.arcq_version().lower()


def test_compute_entry_points_fail_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from arcquant_b200 import agemm
    x = torch.zeros(4, 4096, dtype=torch.bfloat16)
    idx = torch.arange(4096, dtype=torch.int16)
    with pytest.raises(RuntimeError):
        agemm.reorder_quantize_x(x, idx, 0)
    # straight through the C ABI with bogus (non-NULL, aligned) device pointers: status, not a crash or a CPU answer
    L = __import__("arcquant_b200._lib", fromlist=["load"]).load()
    rc = L.arcq_quantize_x(ctypes.c_void_p(0x1000), 4, 4096, ctypes.c_void_p(0x2000), 0, 0, ctypes.c_void_p(0x3000),
                           ctypes.c_void_p(0x4000), None)
    assert rc != 0 and len(L.arcq_last_error()) > 0


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under arcquant_b200/ or agemm/ may reference it."""
    for pkg in ("arcquant_b200", "agemm"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, pkg)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    s = open(os.path.join(dirpath, f)).read()
                    assert not re.search(r"^\s*(from|import)\s+oracle\b", s, flags=re.M), os.path.join(dirpath, f)
                    assert "arcq_oracle" not in s and "libarcq_oracle" not in s
