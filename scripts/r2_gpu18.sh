#!/bin/bash
# GPU call 18 (8 GPUs): BASELINE.json config 5 -- Qwen2.5-32B prefill bs 8 x seq 2048, TP 8: sequence parallel over peer memory vs NCCL all-reduce
mkdir -p gpurun_out/g18
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
A="scripts/prefill_bench.py --model qwen2.5-32b --bsz 8 --seq 2048 --steps 3 --warmup 2"
timeout 400 $T --master-port 29551 $A > gpurun_out/g18/qwen32b_tp8_sp.json 2> gpurun_out/g18/sp.err; echo "sp rc=$?"; cat gpurun_out/g18/qwen32b_tp8_sp.json; grep -v "^\*\|OMP\|^$" gpurun_out/g18/sp.err | tail -5
timeout 400 $T --master-port 29552 $A --no-sp > gpurun_out/g18/qwen32b_tp8_allreduce.json 2> gpurun_out/g18/ar.err; echo "ar rc=$?"; cat gpurun_out/g18/qwen32b_tp8_allreduce.json
