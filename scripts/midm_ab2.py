"""A/B of the kernel choice for 128 < M <= 2048 on the config-2 shapes: single-CTA kernel with 128- or 256-wide tiles
against the cta_group::2 kernel (split-K off).  CUDA graphs, weights rotated (scripts/config2_sweep.py method)."""
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
    for name, (N, K) in shapes.items():
        for M in (192, 256, 384, 512, 768, 1024, 1536, 2048):
            row = {"shape": name, "M": M, "N": N, "K": K}
            for tag, opts in (("default", {}), ("bn128", {"gemm_pair": "0", "gemm_bn": "128", "gemm_splitk": "0"}),
                              ("bn256", {"gemm_pair": "0", "gemm_bn": "256", "gemm_splitk": "0"}), ("pair", {"gemm_pair": "1", "gemm_splitk": "0"})):
                for k, v in opts.items():
                    agemm.set_option(k, v)
                row[tag + "_us"] = round(gemm_point(M, N, K, hbm, tf, calls=6)["us"], 2)
                for k in opts:
                    agemm.set_option(k, "")
            print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
