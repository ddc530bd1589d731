"""profiles/ncu_summary.json from an `ncu --set full` report of scripts/step_prof.py (run here, no GPU needed):
  python scripts/ncu_summary.py gpurun_out/r1_step.ncu-rep [round-tag] [names,in,capture,order]
Also writes profiles/<tag>_step_ncu.txt (the per-launch metric table the numbers come from)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scripts.step_prof import ORDER  # noqa: E402

KEYS = ["Kernel Name", "launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def to_bytes(v, unit):
    f = float(v.replace(",", ""))
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    rep, tag = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "r1")
    global ORDER
    if len(sys.argv) > 3:  # names of the captured launches in capture order (when -s did not land on a round boundary)
        ORDER = sys.argv[3].split(",")
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    assert len(body) == len(ORDER), "expected %d launches, got %d" % (len(ORDER), len(body))
    summ, lines = {}, ["|".join(k for k in KEYS if k in hdr)]
    for name, r in zip(ORDER, body):
        g = lambda k: r[hdr.index(k)]
        u = lambda k: units[hdr.index(k)]
        summ[name] = {
            "kernel": g("Kernel Name")[:80], "ncu_duration_us": float(g("gpu__time_duration.sum")),
            "dram_bytes": to_bytes(g("dram__bytes_read.sum"), u("dram__bytes_read.sum")) + to_bytes(g("dram__bytes_write.sum"), u("dram__bytes_write.sum")),
            "dram_pct_of_peak": float(g("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")),
            "tensor_pipe_active_pct": float(g("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")),
            "issue_active_pct": float(g("smsp__issue_active.avg.pct_of_peak_sustained_active")),
            "registers": int(g("launch__registers_per_thread")), "grid": int(g("launch__grid_size")),
        }
        lines.append(name + "|" + "|".join(r[hdr.index(k)][:70] for k in KEYS if k in hdr))
    summ["_source"] = "ncu --set full --clock-control none of scripts/step_prof.py (last round), report %s" % os.path.basename(rep)
    with open(os.path.join(ROOT, "profiles", "ncu_summary.json"), "w") as f:
        json.dump(summ, f, indent=1)
    with open(os.path.join(ROOT, "profiles", "%s_step_ncu.txt" % tag), "w") as f:
        f.write("\n".join(lines) + "\n")
    print(json.dumps(summ, indent=1))


if __name__ == "__main__":
    main()
