#!/bin/bash
# GPU call 17 (N GPUs): sequence-parallel prefill over peer memory -- check against the all-reduce path, then timings
mkdir -p gpurun_out/g17
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $T --master-port 29541 scripts/prefill_bench.py --model tiny --bsz 2 --seq 512 --steps 2 --warmup 1 --check > gpurun_out/g17/check_tiny.json 2> gpurun_out/g17/check_tiny.err; echo "tiny rc=$?"; cat gpurun_out/g17/check_tiny.json; grep -v "^\*\|OMP" gpurun_out/g17/check_tiny.err | tail -8
M=${2:-qwen2.5-7b}
timeout 300 $T --master-port 29542 scripts/prefill_bench.py --model $M --check > gpurun_out/g17/sp_${M}_tp$N.json 2> gpurun_out/g17/sp.err; echo "sp rc=$?"; cat gpurun_out/g17/sp_${M}_tp$N.json; grep -v "^\*\|OMP" gpurun_out/g17/sp.err | tail -5
timeout 300 $T --master-port 29543 scripts/prefill_bench.py --model $M --no-sp > gpurun_out/g17/ar_${M}_tp$N.json 2> gpurun_out/g17/ar.err; echo "ar rc=$?"; cat gpurun_out/g17/ar_${M}_tp$N.json
timeout 300 $T --master-port 29544 scripts/prefill_bench.py --model $M --no-graph > gpurun_out/g17/sp_nograph_${M}_tp$N.json 2> gpurun_out/g17/spng.err; echo "sp-nograph rc=$?"; cat gpurun_out/g17/sp_nograph_${M}_tp$N.json
