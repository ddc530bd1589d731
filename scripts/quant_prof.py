"""Quantise-kernel timing / profiling driver (GPU box).

  python scripts/quant_prof.py [--shapes M,KQ,KE ...] [--iters 20] [--rms]
Prints one line per shape: us, algorithmic GB/s (SURVEY.md §8d bytes), fraction of MEASURED_PEAKS hbm_gbs.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcquant_b200 import agemm  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", nargs="*", default=["8192,4096,256", "8192,14336,512", "8192,3584,256", "8192,18944,512",
                                                    "8192,28672,1024", "8192,8192,512", "2048,4096,256", "512,4096,256"])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--rms", action="store_true")
    ap.add_argument("--add", action="store_true", help="rmsnorm with the fused residual add (add_rmsnorm_quantize_x) and, beside it, "
                                                     "torch add + rmsnorm_quantize_x")
    ap.add_argument("--identity", action="store_true", help="identity reorder_index: conflict-free gather (upper bound)")
    ap.add_argument("--tile", default="", help='force a tile organisation: "R,threads,S,RT,CTAs/SM" (arcq_set_option quant_tile)')
    ap.add_argument("--no-plans", action="store_true", help="arithmetic chunk placement instead of the conflict-free plans (A/B)")
    args = ap.parse_args()
    peak = 6540.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    dev = torch.device("cuda:0")
    agemm.use_plans(not args.no_plans)
    agemm.set_option("quant_tile", args.tile)
    for sh in args.shapes:
        M, KQ, KE = [int(v) for v in sh.split(",")]
        g = torch.Generator().manual_seed(KQ)
        # several distinct input buffers so consecutive launches do not hit in the 126 MB L2
        nbuf = max(2, int(400e6 // (M * KQ * 2)) + 1)
        xs = [(torch.randn(M, KQ, generator=g) * 50).to(torch.bfloat16).to(dev) for _ in range(min(nbuf, 8))]
        idx = torch.randperm(KQ, generator=g).to(torch.int16).to(dev)
        if args.identity:
            idx = torch.arange(KQ, dtype=torch.int16, device=dev)
        w = torch.ones(KQ, dtype=torch.bfloat16, device=dev)
        rms = args.rms and 2048 <= KQ <= 8192

        res = xs[-1]
        # mode: "" plain, "rms", "add" = fused residual add + rmsnorm, "sep" = torch add, then rmsnorm_quantize_x
        modes = (["add", "sep"] if args.add else ["rms"]) if rms else [""]
        for mode in modes:
            def run(x, mode=mode):
                if mode == "sep":
                    return agemm.rmsnorm_quantize_x(res + x, w, 1e-6, idx, KE)
                if mode == "add":
                    return agemm.add_rmsnorm_quantize_x(x, res, w, 1e-6, idx, KE)
                if mode == "rms":
                    return agemm.rmsnorm_quantize_x(x, w, 1e-6, idx, KE)
                return agemm.reorder_quantize_x(x, idx, KE)
            for i in range(3):
                run(xs[i % len(xs)])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.iters):
                run(xs[i % len(xs)])
            e1.record()
            torch.cuda.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / args.iters
            # algorithmic bytes: the fused add reads a second bf16 tensor and writes the bf16 sum
            by = M * (2 * KQ + (KQ + KE) // 2 + (KQ + KE) // 16) + 2 * KQ + (M * 4 * KQ if mode in ("add", "sep") else 0)
            gbs = by / us / 1e3
            tag = {"": "", "rms": "+rms", "add": "+rms+add(fused)", "sep": "+rms+add(torch add, then rmsnorm_quantize_x)"}[mode]
            print("quant%s M=%d KQ=%d KE=%d: %.1f us  %.0f GB/s  %.3f of measured %.0f" % (tag, M, KQ, KE, us, gbs, gbs / peak, peak), flush=True)


if __name__ == "__main__":
    main()
