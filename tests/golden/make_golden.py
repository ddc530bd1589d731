#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/ -- run in the authoring container only.

Sources (nothing here runs at test time; /root/reference does not exist on the GPU box):
  * oracle/_ref/tables/*.bin   written by oracle/_ref/ref_probe = the reference's vendored CUTLASS v4.3.1
    host converters and CuTe scale-factor layout (oracle/ref_probe.cu, built by oracle/build_ref.sh probe);
  * /root/reference/model/quantize.py imported AS IS (agemm stubbed, its only non-torch import, :7-8):
    fake_reorder_quantize_x / _w outputs on small seeded inputs.

  python tests/golden/make_golden.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
TAB = os.path.join(ROOT, "oracle", "_ref", "tables")


def rd(name, dtype):
    return np.fromfile(os.path.join(TAB, name), dtype=dtype)


def converters():
    out = {
        "e2m1_of_bf16": rd("e2m1_of_bf16.bin", np.uint8),
        "e2m1_boundary_in": rd("e2m1_boundary_in.bin", np.uint32),
        "e2m1_boundary_out": rd("e2m1_boundary_out.bin", np.uint8),
        "e2m1_decode": rd("e2m1_decode.bin", np.float32),
        "ue4m3_in": rd("ue4m3_in.bin", np.uint32),
        "ue4m3_out": rd("ue4m3_out.bin", np.uint8),
        "ue4m3_decode": rd("ue4m3_decode.bin", np.float32),
        "bf16_in": rd("bf16_in.bin", np.uint32),
        "bf16_out": rd("bf16_out.bin", np.uint16),
    }
    np.savez_compressed(os.path.join(HERE, "cutlass_converters.npz"), **out)
    sf = {}
    for rows, K in ((300, 4160), (128, 64), (257, 1088)):
        sf["off_%dx%d" % (rows, K)] = rd("sf_offsets_%dx%d.bin" % (rows, K), np.int32).reshape(rows, K // 16)
        sf["cosize_%dx%d" % (rows, K)] = rd("sf_cosize_%dx%d.bin" % (rows, K), np.int32)
    np.savez_compressed(os.path.join(HERE, "cute_sf_layout.npz"), **sf)


def fakequant():
    sys.modules["agemm"] = types.ModuleType("agemm")  # model/quantize.py:7-8 imports it and never uses it here
    sys.path.insert(0, "/root/reference/model")
    import quantize as refq  # noqa: E402  (the reference module, unmodified)
    from tests._util import make_acts
    out = {}
    for tag, (M, KQ, KE, dist, seed) in {"a": (24, 512, 128, "A", 11), "b": (16, 1024, 64, "B", 45510),
                                          "c": (8, 256, 0, "A", 3)}.items():
        x, _ = make_acts(M, KQ, KE, seed=seed, dist=dist)
        # identity order with the outliers last: the only case where the fake path (primary groups in ORIGINAL
        # channel order, quantize.py:262-268) and the kernels (groups in reordered order) see the same groups --
        # it is also what the reference's own kernels/main.py uses (arange index).
        amax = x.float().abs().amax(0)
        order = torch.argsort(amax, stable=True)
        x = x[:, order].contiguous()
        idx = torch.arange(KQ)
        xf = x.float()
        qx, scale_x, scale = refq.fake_reorder_quantize_x(xf, idx, KE)
        g = torch.Generator().manual_seed(seed + 1)
        w = (torch.randn(32, KQ, generator=g) * 0.02).to(torch.bfloat16).float()
        qw, scale_w, sw = refq.fake_reorder_quantize_w(w, idx, KE)
        out[tag + "_x_bits"] = x.view(torch.int16).numpy().view(np.uint16)
        out[tag + "_KE"] = np.int32(KE)
        out[tag + "_qx"] = qx.numpy().astype(np.float32)
        out[tag + "_scale"] = np.float32(float(scale))
        out[tag + "_w"] = w.numpy()
        out[tag + "_qw"] = qw.numpy().astype(np.float32)
        out[tag + "_sw"] = np.float32(float(sw))
    # a REAL permutation (ascending-absmax argsort of un-sorted channels, utilize.py:42): pins the
    # `topk = reorder_index[-select_num:]` gather of model/quantize.py:237-241,262-267
    for tag, (M, KQ, KE, dist, seed) in {"d": (16, 512, 64, "A", 21), "e": (8, 768, 128, "B", 22)}.items():
        x, idx16 = make_acts(M, KQ, KE, seed=seed, dist=dist)
        idx = idx16.long()
        assert not torch.equal(idx, torch.arange(KQ))
        qx, scale_x, scale = refq.fake_reorder_quantize_x(x.float(), idx, KE)
        g = torch.Generator().manual_seed(seed + 1)
        w = (torch.randn(32, KQ, generator=g) * 0.02).to(torch.bfloat16).float()
        qw, scale_w, sw = refq.fake_reorder_quantize_w(w, idx, KE)
        out[tag + "_x_bits"] = x.view(torch.int16).numpy().view(np.uint16)
        out[tag + "_idx"] = idx.numpy().astype(np.int32)
        out[tag + "_KE"] = np.int32(KE)
        out[tag + "_qx"] = qx.numpy().astype(np.float32)
        out[tag + "_scale"] = np.float32(float(scale))
        out[tag + "_w"] = w.numpy()
        out[tag + "_qw"] = qw.numpy().astype(np.float32)
        out[tag + "_sw"] = np.float32(float(sw))
    np.savez_compressed(os.path.join(HERE, "fakequant_reference.npz"), **out)


if __name__ == "__main__":
    converters()
    fakequant()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
