"""Which side limits the GEMM mainloop?  Times the plain kernel with loads / MMAs / UTCCP disabled."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arcquant_b200 import agemm
from scripts.gemm_sweep import rand_operands
dev = torch.device("cuda:0")
os.environ["ARCQ_GEMM_CLUSTER"] = "0"; os.environ["ARCQ_GEMM_PAIR"] = "1"
for (M, N, K) in [(8192, 14336, 4352), (8192, 4096, 14848)]:
    a, b, sfa, sfb = rand_operands(M, N, K, dev)
    y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
    for bn in ("256",):
        os.environ["ARCQ_GEMM_BN"] = bn
        for dbg, name in [(0, "full"), (1, "no-loads"), (6, "loads only"), (5, "mma only"), (7, "nothing"), (8, "full, no stores"), (14, "loads only, no stores"), (13, "mma only, no stores")]:
            os.environ["ARCQ_GEMM_DEBUG"] = str(dbg)
            for _ in range(3):
                agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
            e1.record(); torch.cuda.synchronize()
            us = 100 * e0.elapsed_time(e1)
            print("M=%d N=%d K=%d BN=%s %-20s %.1f us (%.0f TFLOP/s-equivalent)" % (M, N, K, bn, name, us, 2.0*M*N*K/us/1e6), flush=True)
