"""Launch each distinct kernel shape of the (default, fused) bench step once per round, in a fixed order, for
`ncu -k regex:'quant|gemm'`.

  python scripts/step_prof.py [--rounds 3]     # ncu: -s 6*(rounds-1) -c 6 captures the last round
Order per round (M=8192): quant KQ4096/KE256, gemm q+k+v N6144 K4352 (per-column scales), gemm o N4096 K4352,
                          gemm gate/up + SwiGLU N28672 K4352, quant KQ14336/KE512, gemm down N4096 K14848.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcquant_b200 import agemm  # noqa: E402
from tests._util import make_acts  # noqa: E402

ORDER = ["quant_KQ4096", "gemm_N6144_K4352", "gemm_N4096_K4352", "gemm_N28672_K4352", "quant_KQ14336", "gemm_N4096_K14848"]


def qw(N, KQ, KE, idx, g, dev):
    w = (torch.randn(N, KQ, generator=g) * 0.02).to(torch.bfloat16)
    sw = float(w.max().float() / 2688.0)
    q, sf = agemm.reorder_quantize_w((w / sw).to(torch.bfloat16).to(dev), idx, KE)
    return q, sf, sw


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--tokens", type=int, default=8192)
    args = ap.parse_args()
    M, dev = args.tokens, torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    sx = 8.0 / 2688.0
    xs, idx = make_acts(256, 4096, 256, seed=4096)
    x1, idx1 = xs.repeat(M // 256, 1).contiguous().to(dev), idx.to(dev)
    qkv = agemm.fuse_weights([qw(n, 4096, 256, idx1, g, dev) for n in (4096, 1024, 1024)])
    o = qw(4096, 4096, 256, idx1, g, dev)
    gu = agemm.interleave_gate_up(qw(14336, 4096, 256, idx1, g, dev), qw(14336, 4096, 256, idx1, g, dev))
    xs, idx = make_acts(256, 14336, 512, seed=14336)
    x2, idx2 = xs.repeat(M // 256, 1).contiguous().to(dev), idx.to(dev)
    down = qw(4096, 14336, 512, idx2, g, dev)
    torch.cuda.synchronize()
    for _ in range(args.rounds):
        qx, sfx = agemm.reorder_quantize_x(x1, idx1, 256)
        agemm.matmul(qx, qkv[0], sfx, qkv[1], sx, col_scale=qkv[2])
        agemm.matmul(qx, o[0], sfx, o[1], sx * o[2])
        agemm.matmul(qx, gu[0], sfx, gu[1], sx, col_scale=gu[2], epilogue=agemm.EPILOGUE_SWIGLU)
        qx2, sfx2 = agemm.reorder_quantize_x(x2, idx2, 512)
        agemm.matmul(qx2, down[0], sfx2, down[1], sx * down[2])
        torch.cuda.synchronize()
    print("order:", ",".join(ORDER))


if __name__ == "__main__":
    main()
