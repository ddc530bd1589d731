"""Launch each distinct kernel shape of the bench step once per round, in a fixed order, for `ncu -k regex:'quant|gemm'`.

  python scripts/step_prof.py [--rounds 3]     # ncu: -s 6*(rounds-1) -c 6 captures the last round
Order per round (M=8192): quant KQ4096/KE256, gemm N4096 K4352, gemm N1024 K4352, gemm N14336 K4352,
                          quant KQ14336/KE512, gemm N4096 K14848.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcquant_b200 import agemm  # noqa: E402
from tests._util import make_acts  # noqa: E402

ORDER = ["quant_KQ4096", "gemm_N4096_K4352", "gemm_N1024_K4352", "gemm_N14336_K4352", "quant_KQ14336", "gemm_N4096_K14848"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--tokens", type=int, default=8192)
    args = ap.parse_args()
    M, dev = args.tokens, torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    st = {}
    for KQ, KE, Ns in ((4096, 256, (4096, 1024, 14336)), (14336, 512, (4096,))):
        xs, idx = make_acts(256, KQ, KE, seed=KQ)
        x = xs.repeat(M // 256, 1).contiguous().to(dev)
        idx = idx.to(dev)
        ws = []
        for N in Ns:
            w = (torch.randn(N, KQ, generator=g) * 50).to(torch.bfloat16).to(dev)
            ws.append(agemm.reorder_quantize_w(w, idx, KE))
        st[KQ] = (x, idx, KE, ws)
    torch.cuda.synchronize()
    for _ in range(args.rounds):
        for KQ in (4096, 14336):
            x, idx, KE, ws = st[KQ]
            qx, sfx = agemm.reorder_quantize_x(x, idx, KE)
            for qw, sfw in ws:
                agemm.matmul(qx, qw, sfx, sfw, 1e-4)
        torch.cuda.synchronize()
    print("order:", ",".join(ORDER))


if __name__ == "__main__":
    main()
