"""A/B of the split-K kernel over several 128-row tiles (128 < M <= 1024) against the one-CTA-per-tile dispatch, on the
Llama-3-8B shapes of BASELINE.json config 2.  CUDA graphs, weights rotated (scripts/config2_sweep.py method)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from config2_sweep import agemm, gemm_point, peaks  # noqa: E402


def main():
    hbm, tf = peaks()
    shapes = {"q/o": (4096, 4352), "k/v": (1024, 4352), "qkv": (6144, 4352), "gate/up": (14336, 4352), "down": (4096, 14848)}
    out = []
    for name, (N, K) in shapes.items():
        for M in (64, 128, 192, 256, 384, 512, 768, 1024):
            row = {"shape": name, "M": M, "N": N, "K": K}
            for tag, opt in (("default", ""), ("no_splitk", "0"), ("s2", "2"), ("s4", "4"), ("s8", "8")):
                agemm.set_option("gemm_splitk", opt)
                row[tag + "_us"] = round(gemm_point(M, N, K, hbm, tf)["us"], 2)
            agemm.set_option("gemm_splitk", "")
            out.append(row)
            print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
