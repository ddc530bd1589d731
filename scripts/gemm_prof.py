import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arcquant_b200 import agemm
from scripts.gemm_sweep import rand_operands
dev = torch.device("cuda:0")
os.environ["ARCQ_GEMM_CLUSTER"] = "0"; os.environ["ARCQ_GEMM_PAIR"] = "1"
for (M, N, K) in [(8192, 14336, 4352), (8192, 4096, 14848)]:
    a, b, sfa, sfb = rand_operands(M, N, K, dev)
    y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
    for _ in range(3):
        agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
    torch.cuda.synchronize()
    for dbg in ("0",):
        os.environ["ARCQ_GEMM_DEBUG"] = dbg
        os.environ["ARCQ_GEMM_PROF"] = "1"
        print("M=%d N=%d K=%d debug=%s" % (M, N, K, dbg), file=sys.stderr, flush=True)
        agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
        agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
        del os.environ["ARCQ_GEMM_PROF"]
