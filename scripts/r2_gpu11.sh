#!/bin/bash
# GPU call 11 (8 GPUs): peer-memory exchange at 8 ranks (llama3-8b layer = the driver's scaling workload; llama3-70b = config 4)
mkdir -p gpurun_out
N=${1:-8}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
B="bench.py --gpus $N --no-cpu-baseline --no-prefill --e2e-steps 2"
timeout 240 $T --master-port 29521 $B --steps 30 --warmup 5 > gpurun_out/g11_bench_tp${N}_peer.json 2> gpurun_out/g11_bench_tp${N}_peer.err; echo "peer rc=$?"
timeout 240 $T --master-port 29522 $B --steps 20 --warmup 5 --workload llama3-70b > gpurun_out/g11_bench_tp${N}_peer_70b.json 2> gpurun_out/g11_bench_tp${N}_peer_70b.err; echo "70b rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/g11_bench_*.json")):
    try:
        o=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, o["ms_per_step"], o["value"], o.get("exchange"), o.get("exchange_note"), o["parity_check"])
        print("   ", [(k["kernel"], k.get("launch") or k.get("input_of") or k.get("of"), round(k["us"],1)) for k in o["kernels"]])
    except Exception as e:
        print(f, "ERR", e)
PY
tail -5 gpurun_out/g11_bench_tp${N}_peer.err
