#!/bin/bash
# GPU call 10 (2 GPUs): peer-memory exchange -- kernel tests, 2-rank bench A/B (peer vs NCCL), 2-rank parity runs
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/g10_topo.txt 2>&1
#timeout 400 python -m pytest tests/test_gpu_peer.py -x -q > gpurun_out/g10_peer_tests.log 2>&1; echo "peer tests rc=$?"
tail -3 gpurun_out/g10_peer_tests.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
B="bench.py --gpus 2 --no-cpu-baseline --no-prefill --e2e-steps 2"
timeout 300 $T --master-port 29511 $B --steps 30 --warmup 5 > gpurun_out/g10_bench_tp2_peer.json 2> gpurun_out/g10_bench_tp2_peer.err; echo "peer rc=$?"
timeout 300 $T --master-port 29512 $B --steps 30 --warmup 5 --no-peer > gpurun_out/g10_bench_tp2_nccl.json 2> gpurun_out/g10_bench_tp2_nccl.err; echo "nccl rc=$?"
timeout 300 $T --master-port 29513 $B --steps 30 --warmup 5 --graph > gpurun_out/g10_bench_tp2_peer_graph.json 2> gpurun_out/g10_bench_tp2_peer_graph.err; echo "graph rc=$?"
timeout 300 $T --master-port 29514 $B --steps 20 --warmup 5 --workload llama3-70b > gpurun_out/g10_bench_tp2_peer_70b.json 2> gpurun_out/g10_bench_tp2_peer_70b.err; echo "70b rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/g10_bench_*.json")):
    try:
        o=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, o["ms_per_step"], o["value"], o.get("exchange"), o.get("exchange_note"), o["parity_check"])
        print("   ", [(k["kernel"], k.get("launch") or k.get("input_of"), round(k["us"],1)) for k in o["kernels"]])
    except Exception as e:
        print(f, "ERR", e)
PY
tail -5 gpurun_out/g10_bench_tp2_peer.err
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -k "args1 or args6 or one_process" > gpurun_out/g10_multi_tests.log 2>&1; echo "multi tests rc=$?"
tail -5 gpurun_out/g10_multi_tests.log
