"""Sustained GEMM loop with nvidia-smi clock/power sampling (is the tensor pipe power/clock limited?)."""
import os, sys, subprocess, time, statistics, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arcquant_b200 import agemm
from scripts.gemm_sweep import rand_operands
dev = torch.device("cuda:0")
os.environ["ARCQ_GEMM_CLUSTER"] = "0"
M, N, K = 8192, 14336, 4352
a, b, sfa, sfb = rand_operands(M, N, K, dev)
y = torch.empty((M, N), dtype=torch.bfloat16, device=dev)
for dbg in ("0", "5"):
    os.environ["ARCQ_GEMM_DEBUG"] = dbg
    for _ in range(3):
        agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
    torch.cuda.synchronize()
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown,temperature.gpu",
                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
    time.sleep(0.3)
    res = []
    for rep in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(1500):
            agemm.matmul(a, b, sfa, sfb, 1e-3, out=y)
        e1.record(); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / 1500 * 1e3)
    time.sleep(0.1)
    p.terminate(); out = p.communicate()[0]
    rows = [l.split(",") for l in out.strip().splitlines()]
    clk = [float(r[0]) for r in rows if len(r) >= 6]
    pw = [float(r[1]) for r in rows if len(r) >= 6]
    cap = sum(1 for r in rows if len(r) >= 6 and "Active" in r[2] and "Not" not in r[2])
    print("debug=%s us per GEMM per 1500-iter block: %s" % (dbg, ["%.1f" % v for v in res]))
    print("  clocks MHz: first %s ... median %.0f min %.0f max %.0f ; power W median %.0f max %.0f ; sw_power_cap active in %d/%d samples" % (
        clk[:6], statistics.median(clk), min(clk), max(clk), statistics.median(pw), max(pw), cap, len(rows)))
    print("  clock trace:", clk[::4])
