#!/bin/bash
# GPU call 14 (N GPUs): P2P store bandwidth probe (first two GPUs) + N-rank bench on the peer-memory exchange
mkdir -p gpurun_out
N=${1:-4}
timeout 120 python scripts/peer_bw_probe.py > gpurun_out/g14_peer_bw.json 2> gpurun_out/g14_peer_bw.err; cat gpurun_out/g14_peer_bw.json; tail -3 gpurun_out/g14_peer_bw.err
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
B="bench.py --gpus $N --no-cpu-baseline --no-prefill --e2e-steps 2"
timeout 240 $T --master-port 29531 $B --steps 30 --warmup 5 > gpurun_out/g14_bench_tp${N}_peer.json 2> gpurun_out/g14_bench_tp${N}_peer.err; echo "peer rc=$?"
timeout 240 $T --master-port 29532 $B --steps 30 --warmup 5 --no-peer > gpurun_out/g14_bench_tp${N}_nccl.json 2> gpurun_out/g14_bench_tp${N}_nccl.err; echo "nccl rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/g14_bench_*.json")):
    try:
        o=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, o["ms_per_step"], o["value"], o.get("exchange"), o.get("exchange_note"), o["parity_check"]["ok"])
        print("   ", [(k["kernel"], k.get("launch") or k.get("input_of") or k.get("of"), round(k["us"],1)) for k in o["kernels"]])
    except Exception as e:
        print(f, "ERR", e)
PY
