# round 2, GPU call 5 (2 GPUs, strictly bounded): exit-hang fix, OWN2 + silu-table validation, TP2 lines
set -x
mkdir -p gpurun_out/g5
timeout 600 python -m pytest tests -q -m gpu --maxfail=20 -p no:cacheprovider --timeout 300 > gpurun_out/g5/tests_gpu.log 2>&1; tail -15 gpurun_out/g5/tests_gpu.log | cut -c1-300
run() { tag=$1; shift; timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline --no-prefill "$@" > gpurun_out/g5/bench_$tag.json 2> gpurun_out/g5/bench_$tag.err; echo "rc=$?"; tail -c 400 gpurun_out/g5/bench_$tag.json; tail -2 gpurun_out/g5/bench_$tag.err | cut -c1-300; }
run tp2
run tp2_graph --graph
run tp2_ar --parallelism tp-ar
run tp2_70b --workload llama3-70b
timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/g5/bench_n1.json 2> gpurun_out/g5/bench_n1.err; tail -c 400 gpurun_out/g5/bench_n1.json
timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-prefill --swiglu > gpurun_out/g5/bench_n1_swiglu.json 2> gpurun_out/g5/bench_n1_swiglu.err; tail -c 300 gpurun_out/g5/bench_n1_swiglu.json
timeout 120 python scripts/quant_prof.py --shapes 8192,4096,256 8192,14336,512 8192,8192,512 32768,14336,512 > gpurun_out/g5/quant_own2.log 2>&1; cat gpurun_out/g5/quant_own2.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/prefill_bench.py --model qwen2.5-7b --bsz 4 --seq 1024 > gpurun_out/g5/prefill_qwen7b_tp2.json 2> gpurun_out/g5/prefill_tp2.err; cat gpurun_out/g5/prefill_qwen7b_tp2.json; tail -2 gpurun_out/g5/prefill_tp2.err | cut -c1-300
