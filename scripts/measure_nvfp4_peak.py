"""Measured dense-NVFP4 denominator + the "numbers to beat" of BASELINE.md §1, taken on the box (one GPU):

  1. cuBLASLt NVFP4 through torch._scaled_mm (float4_e2m1fn_x2 operands, swizzled E4M3 block scales -- the very
     buffers our quantise kernel writes) at 8192^3 and at the bench shapes: burst (best of 10) and sustained (back to back
     for ~3 s); our GEMM on the same operands beside it, and the two results compared (independent parity evidence);
  2. the reference's OWN quantise kernels (oracle/_ref/libarcq_refkernels.so = kernels/src/{reorder,rmsnorm}.cu compiled
     unchanged for sm_100a) timed at the bench shapes next to ours (the oracle library is loaded here, in this untimed
     measurement script, never by the product).

  python scripts/measure_nvfp4_peak.py > profiles/r2_nvfp4_peak.json
"""
import ctypes
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcquant_b200 import agemm  # noqa: E402

dev = torch.device("cuda:0")
L2_FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def time_fn(fn, iters=10, flush=True):
    """best and median of `iters` single launches, each after an L2 flush (CUDA events)."""
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            L2_FLUSH.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[0], ts[len(ts) // 2]


def sustained(fn, seconds=3.0):
    fn()
    torch.cuda.synchronize()
    n, t0 = 0, time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.perf_counter() - t0 < seconds:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def rand_operand(rows, K, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    q = torch.randint(0, 256, (rows, K // 2), dtype=torch.uint8, device=dev, generator=g)
    sf = torch.randint(0x30, 0x40, (agemm.sf_bytes(rows, K),), dtype=torch.uint8, device=dev, generator=g)
    return q, sf


def scaled_mm_callable(qa, sfa, qb, sfb, M, N, K):
    """cuBLASLt NVFP4 on OUR buffers.  Returns (callable, how) or (None, error)."""
    f4 = getattr(torch, "float4_e2m1fn_x2", None)
    if f4 is None:
        return None, "torch has no float4_e2m1fn_x2"
    a = qa.view(f4)
    b = qb.view(f4).t()
    na = (M + 127) // 128 * 128 * K // 16
    nb = (N + 127) // 128 * 128 * K // 16
    sa = sfa[:na].view(torch.float8_e4m3fn)
    sb = sfb[:nb].view(torch.float8_e4m3fn)
    errs = []
    for how, mk in (("torch._scaled_mm(flat scales)", lambda: (lambda: torch._scaled_mm(a, b, sa, sb, out_dtype=torch.bfloat16))),
                    ("torch._scaled_mm(2-D scales)", lambda: (lambda: torch._scaled_mm(a, b, sa.view(-1, K // 16), sb.view(-1, K // 16), out_dtype=torch.bfloat16)))):
        try:
            fn = mk()
            fn()
            torch.cuda.synchronize()
            return fn, how
        except Exception as e:  # noqa: BLE001
            errs.append("%s: %s" % (how, str(e).splitlines()[0][:200]))
    return None, " | ".join(errs)


def load_refgemm():
    """oracle/_ref/libarcq_refgemm.so: the reference's CUTLASS GEMM retargeted Sm120 -> Sm100 (oracle/ref_cutlass_gemm.cu)."""
    so = os.path.join(ROOT, "oracle", "_ref", "libarcq_refgemm.so")
    if not os.path.exists(so):
        return None
    L = ctypes.CDLL(so)
    vp, i = ctypes.c_void_p, ctypes.c_int
    L.arcq_refgemm.argtypes = [i, vp, vp, vp, vp, i, i, i, ctypes.c_float, vp, vp, ctypes.c_longlong, vp]
    L.arcq_refgemm.restype = i
    return L


REFGEMM_VARIANTS = {0: "256x256x256 cluster 2x4 (CUTLASS example 72a)", 1: "256x256x256 cluster 2x1", 2: "128x128x256 cluster 1x1 (nvfp4.cu:32-33 tile)"}


def gemm_section():
    out = []
    RG = load_refgemm()
    ws = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
    shapes = [(8192, 8192, 8192), (8192, 14336, 4352), (8192, 4096, 14848), (8192, 4096, 4352), (8192, 1024, 4352), (8192, 6144, 4352),
              (8192, 28672, 4352)]
    for M, N, K in shapes:
        qa, sfa = rand_operand(M, K, 1)
        qb, sfb = rand_operand(N, K, 2)
        flops = 2.0 * M * N * K
        rec = {"M": M, "N": N, "K": K}
        y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
        ours = lambda: agemm.matmul(qa, qb, sfa, sfb, 1e-3, out=y)  # noqa: E731
        best, med = time_fn(ours)
        rec["ours_us_best"], rec["ours_us_median"] = best, med
        rec["ours_tflops_burst"] = flops / best / 1e6
        rec["ours_tflops_sustained"] = flops / sustained(ours) / 1e6
        fn, how = scaled_mm_callable(qa, sfa, qb, sfb, M, N, K)
        rec["cublaslt_how"] = how
        if fn is not None:
            best, med = time_fn(fn)
            rec["cublaslt_us_best"], rec["cublaslt_us_median"] = best, med
            rec["cublaslt_tflops_burst"] = flops / best / 1e6
            rec["cublaslt_tflops_sustained"] = flops / sustained(fn) / 1e6
            ref = fn().float() * 1e-3
            got = y.float()
            rec["ours_vs_cublaslt_rel_fro"] = float((got - ref).norm() / ref.norm())
            rec["ours_over_cublaslt_burst"] = rec["ours_tflops_burst"] / rec["cublaslt_tflops_burst"]
        if RG is not None:
            st = torch.cuda.current_stream().cuda_stream
            for v, what in REFGEMM_VARIANTS.items():
                yr = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
                call = lambda: RG.arcq_refgemm(v, qa.data_ptr(), qb.data_ptr(), sfa.data_ptr(), sfb.data_ptr(), M, N, K, 1e-3, yr.data_ptr(),  # noqa: E731
                                               ws.data_ptr(), ws.numel(), st)
                rc = call()
                torch.cuda.synchronize()
                if rc != 0:
                    rec["ref_cutlass_v%d" % v] = {"config": what, "error": rc}
                    continue
                best, med = time_fn(call)
                rec["ref_cutlass_v%d" % v] = {"config": what, "us_best": best, "tflops_burst": flops / best / 1e6,
                                              "ours_vs_ref_rel_fro": float((y.float() - yr.float()).norm() / yr.float().norm()),
                                              "ours_speedup": best / rec["ours_us_best"]}
        out.append(rec)
        del qa, sfa, qb, sfb, y
    return out


def quant_section():
    so = os.path.join(ROOT, "oracle", "_ref", "libarcq_refkernels.so")
    R = None
    if os.path.exists(so):
        R = ctypes.CDLL(so)
        vp, i = ctypes.c_void_p, ctypes.c_int
        R.arcq_ref_quantize_x.argtypes = [vp, i, i, vp, i, vp, vp]
        R.arcq_ref_rmsnorm_quantize_x.argtypes = [vp, vp, ctypes.c_float, i, i, vp, i, vp, vp]
    out = []
    for M, KQ, KE in [(8192, 4096, 256), (8192, 14336, 512), (8192, 8192, 512), (8192, 3584, 256), (32768, 4096, 256), (2048, 4096, 256),
                      (512, 4096, 256)]:
        g = torch.Generator(device=dev).manual_seed(KQ)
        x = (torch.randn(M, KQ, device=dev, generator=g) * 40).to(torch.bfloat16)
        idx = torch.randperm(KQ, device=dev, generator=g).to(torch.int16)
        by = M * (2 * KQ + (KQ + KE) // 2 + (KQ + KE) // 16) + 2 * KQ
        rec = {"M": M, "KQ": KQ, "KE": KE, "algorithmic_bytes": by}
        best, med = time_fn(lambda: agemm.reorder_quantize_x(x, idx, KE))
        rec["ours_us_best"], rec["ours_us_median"], rec["ours_GBps"] = best, med, by / best / 1e3
        best, med = time_fn(lambda: agemm.reorder_quantize_x_scaled(x, idx, KE))
        rec["ours_scaled_us_best"] = best
        amax = agemm.absmax(x)
        best, med = time_fn(lambda: agemm.reorder_quantize_x_scaled(x, idx, KE, amax=amax))
        rec["ours_scaled_carried_amax_us_best"] = best
        best, med = time_fn(lambda: agemm.absmax(x))
        rec["absmax_us_best"] = best

        def torch_prescale():
            scale = torch.max(x.abs()).float() / (448.0 * 6.0)
            return agemm.reorder_quantize_x(x / scale, idx, KE)
        best, med = time_fn(torch_prescale)
        rec["torch_prescale_then_quantise_us_best"] = best
        if KQ in (2048, 3072, 3584, 4096, 5120, 8192):
            w = torch.ones(KQ, dtype=torch.bfloat16, device=dev)
            best, med = time_fn(lambda: agemm.rmsnorm_quantize_x(x, w, 1e-6, idx, KE))
            rec["ours_rmsnorm_us_best"] = best
        if R is not None:
            K = KQ + KE
            q = torch.empty((M, K // 2), dtype=torch.uint8, device=dev)
            sf = torch.empty(((M // 128 + 1) * 128 * K // 16,), dtype=torch.uint8, device=dev)
            # the reference launches on the legacy default stream: time it with events on that stream
            with torch.cuda.stream(torch.cuda.default_stream()):
                best, med = time_fn(lambda: R.arcq_ref_quantize_x(x.data_ptr(), M, KQ, idx.data_ptr(), KE, q.data_ptr(), sf.data_ptr()))
                rec["ref_kernel_us_best"], rec["ref_kernel_GBps"] = best, by / best / 1e3
                rec["speedup_vs_ref_kernel"] = rec["ref_kernel_us_best"] / rec["ours_us_best"]
                if "ours_rmsnorm_us_best" in rec:
                    best, med = time_fn(lambda: R.arcq_ref_rmsnorm_quantize_x(x.data_ptr(), w.data_ptr(), 1e-6, M, KQ, idx.data_ptr(), KE,
                                                                              q.data_ptr(), sf.data_ptr()))
                    rec["ref_rmsnorm_kernel_us_best"] = best
        out.append(rec)
    return out


def main():
    res = {"gpu": torch.cuda.get_device_name(0), "torch": torch.__version__, "backend": agemm.BACKEND,
           "method": "CUDA events around single launches after a 256 MB L2 flush, best of 10 (burst); back-to-back launches for 3 s (sustained)"}
    res["gemm"] = gemm_section()
    res["quant"] = quant_section()
    g8 = [r for r in res["gemm"] if (r["M"], r["N"], r["K"]) == (8192, 8192, 8192)]
    if g8 and "cublaslt_tflops_burst" in g8[0]:
        res["nvfp4_peak_tflops_burst"] = max(r.get("cublaslt_tflops_burst", 0) for r in res["gemm"])
        res["nvfp4_peak_tflops_sustained"] = max(r.get("cublaslt_tflops_sustained", 0) for r in res["gemm"])
        res["nvfp4_peak_source"] = "cuBLASLt via torch._scaled_mm, best over the listed shapes"
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
