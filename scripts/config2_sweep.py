"""BASELINE.json config 2 as written: Llama-3-8B linear shapes, quantise + GEMM, M = 1 .. 8192, on one B200.

For every point the kernel is captured ten times in a CUDA graph (host time out of the number; the loop of
/root/reference/kernels/bench.py:9-43 times matmul only, in a Python loop) and the graph is replayed five times.  Weights
rotate over enough copies to exceed the 126 MB L2, so the small-M points really stream their weights from HBM.
Roofline per point: t_min = max(FLOPs / measured NVFP4 peak, algorithmic bytes / measured HBM copy
bandwidth); frac = t_min / t.  Sweeps: M x the four GEMM shapes at the default select_num; KE in {0, 64, 256, 512, 1024} and
the 0.4375 KQ stress of kernels/benchmark/bench_nvfp4.cu:23-25 at M = 64 and M = 8192; the quantise kernel over M.

  python scripts/config2_sweep.py > profiles/r2_sweep.json
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcquant_b200 import agemm  # noqa: E402

dev = torch.device("cuda:0")


def peaks():
    hbm, tf = 6650.0, 6562.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        hbm, tf = d["hbm_gbs"], 4.0 * d["bf16_tflops"]
    p = os.path.join(ROOT, "profiles", "r2_nvfp4_peak.json")
    if os.path.exists(p):
        tf = json.load(open(p)).get("nvfp4_peak_tflops_burst", tf)
    return hbm, tf


def graph_time(fns, reps=5):
    """us per call: every callable of `fns` once per graph, graph replayed `reps` times, best replay."""
    for f in fns[:2]:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for f in fns:
                f()
    torch.cuda.current_stream().wait_stream(s)
    g.replay()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / len(fns))
    return best


def rand_operand(rows, K, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    q = torch.randint(0, 256, (rows, K // 2), dtype=torch.uint8, device=dev, generator=g)
    sf = torch.randint(0x30, 0x40, (agemm.sf_bytes(rows, K),), dtype=torch.uint8, device=dev, generator=g)
    return q, sf


def gemm_point(M, N, K, hbm, tf, calls=10):
    wbytes = 0.5625 * K * N
    ncopies = max(1, min(calls, int(300e6 // wbytes) + 1))
    ws = [rand_operand(N, K, 100 + i) for i in range(ncopies)]
    qa, sfa = rand_operand(M, K, 1)
    y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
    fns = [(lambda w=ws[i % ncopies]: agemm.matmul(qa, w[0], sfa, w[1], 1e-3, out=y)) for i in range(calls)]
    us = graph_time(fns)
    flops = 2.0 * M * N * K
    by = 0.5625 * K * (M + N) + 2.0 * M * N
    t_min = max(flops / (tf * 1e12), by / (hbm * 1e9)) * 1e6
    return {"M": M, "N": N, "K": K, "us": us, "TFLOPs": flops / us / 1e6, "GBps": by / us / 1e3, "bound": "tensor" if flops / (tf * 1e12) > by / (hbm * 1e9) else "hbm",
            "roofline_us": t_min, "frac": t_min / us}


def quant_point(M, KQ, KE, hbm, calls=10):
    xb = M * KQ * 2
    ncopies = max(1, min(calls, int(300e6 // xb) + 1))
    g = torch.Generator(device=dev).manual_seed(KQ)
    xs = [(torch.randn(M, KQ, device=dev, generator=g) * 40).to(torch.bfloat16) for _ in range(ncopies)]
    idx = torch.randperm(KQ, device=dev, generator=g).to(torch.int16)
    agemm.reorder_quantize_x(xs[0], idx, KE)   # builds the placement plan outside the capture
    fns = [(lambda x=xs[i % ncopies]: agemm.reorder_quantize_x(x, idx, KE)) for i in range(calls)]
    us = graph_time(fns)
    by = M * (2 * KQ + (KQ + KE) // 2 + (KQ + KE) // 16) + 2 * KQ
    return {"M": M, "KQ": KQ, "KE": KE, "us": us, "GBps": by / us / 1e3, "roofline_us": by / (hbm * 1e9) * 1e6, "frac": by / us / 1e3 / hbm}


def main():
    hbm, tf = peaks()
    Ms = [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192]
    shapes = {"q/o (4096, 4096+KE)": (4096, 4096, 256), "k/v (1024, 4096+KE)": (1024, 4096, 256), "qkv fused (6144, 4096+KE)": (6144, 4096, 256),
              "gate/up (14336, 4096+KE)": (14336, 4096, 256), "down (4096, 14336+KE)": (4096, 14336, 512)}
    res = {"gpu": torch.cuda.get_device_name(0), "peaks": {"hbm_gbs": hbm, "nvfp4_tflops": tf}, "binding": agemm.BACKEND,
           "method": "10 calls per CUDA graph, best of 5 replays, weights rotated over > 300 MB of copies", "gemm_M_sweep": {}, "gemm_KE_sweep": [],
           "quant_M_sweep": []}
    for name, (N, KQ, KE) in shapes.items():
        res["gemm_M_sweep"][name] = [gemm_point(M, N, KQ + KE, hbm, tf) for M in Ms]
        torch.cuda.empty_cache()
    for M in (64, 8192):
        for N, KQ in ((4096, 4096), (14336, 4096), (4096, 14336)):
            for KE in (0, 64, 256, 512, 1024, int(0.4375 * KQ)):
                r = gemm_point(M, N, KQ + KE, hbm, tf)
                r["KQ"], r["KE"] = KQ, KE
                res["gemm_KE_sweep"].append(r)
        torch.cuda.empty_cache()
    for KQ, KE in ((4096, 256), (14336, 512)):
        for M in Ms[4:]:
            res["quant_M_sweep"].append(quant_point(M, KQ, KE, hbm))
        torch.cuda.empty_cache()
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
