"""GEMM experiment driver (GPU box): times arcq_gemm_nvfp4 variants and checks them against the plain kernel.

  python scripts/gemm_sweep.py [--shapes M,N,K ...] [--clusters 0,21,22,41,42,24] [--iters 20]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arcquant_b200 import agemm  # noqa: E402


def rand_operands(M, N, K, dev):
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    a = torch.randint(0, 256, (M, K // 2), dtype=torch.uint8, generator=g).to(dev)
    b = torch.randint(0, 256, (N, K // 2), dtype=torch.uint8, generator=g).to(dev)
    # scales in [2^-3, 2^2): bytes 0x20..0x47
    sfa = torch.randint(0x20, 0x48, (agemm.sf_bytes(M, K),), dtype=torch.uint8, generator=g).to(dev)
    sfb = torch.randint(0x20, 0x48, (agemm.sf_bytes(N, K),), dtype=torch.uint8, generator=g).to(dev)
    return a, b, sfa, sfb


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", nargs="*", default=["8192,14336,4352", "8192,4096,4352", "8192,4096,14848",
                                                    "8192,1024,4352", "1000,1000,4160"])
    ap.add_argument("--clusters", default="0,21,22,41,42,24")
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    for sh in args.shapes:
        M, N, K = [int(v) for v in sh.split(",")]
        a, b, sfa, sfb = rand_operands(M, N, K, dev)
        os.environ["ARCQ_GEMM_CLUSTER"] = "0"
        os.environ["ARCQ_GEMM_PAIR"] = "0"
        os.environ["ARCQ_GEMM_BN"] = "256"
        base = agemm.matmul(a, b, sfa, sfb, 1e-3)
        torch.cuda.synchronize()
        for cs in args.clusters.split(","):
            os.environ["ARCQ_GEMM_PAIR"] = "1" if cs == "p" else "0"
            os.environ["ARCQ_GEMM_CLUSTER"] = "0" if cs == "p" else cs
            y = torch.empty_like(base)
            for _ in range(3):
                agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
            e1.record()
            torch.cuda.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / args.iters
            same = bool(torch.equal(y, base))
            nbad = int((y != base).sum()) if not same else 0
            print("M=%d N=%d K=%d cluster=%s: %.1f us  %.0f TFLOP/s  identical=%s bad=%d" % (
                M, N, K, cs, us, 2.0 * M * N * K / us / 1e6, same, nbad), flush=True)


if __name__ == "__main__":
    main()
