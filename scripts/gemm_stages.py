import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arcquant_b200 import agemm
from scripts.gemm_sweep import rand_operands
dev = torch.device("cuda:0")
os.environ["ARCQ_GEMM_CLUSTER"] = "0"; os.environ["ARCQ_GEMM_BN"] = "256"
for (M, N, K) in [(8192, 14336, 4352), (8192, 4096, 14848), (8192, 4096, 4352), (8192, 1024, 4352)]:
    a, b, sfa, sfb = rand_operands(M, N, K, dev)
    ys = {}
    for st in ("3", "4"):
        os.environ["ARCQ_GEMM_STAGES"] = st
        y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
        for _ in range(3):
            agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
        e1.record(); torch.cuda.synchronize()
        us = 50 * e0.elapsed_time(e1)
        ys[st] = y
        print("M=%d N=%d K=%d stages=%s %.1f us %.0f TFLOP/s" % (M, N, K, st, us, 2.0*M*N*K/us/1e6), flush=True)
    print("  identical:", bool(torch.equal(ys["3"], ys["4"])))
