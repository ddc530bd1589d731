# round 2, GPU call 4 (2 GPUs): multi-GPU parity tests, TP2 bench variants, TP2 prefill
set -x
mkdir -p gpurun_out/g4
nvidia-smi -L
timeout 1500 python -m pytest tests/test_gpu_multi.py -q -m gpu -p no:cacheprovider --timeout 900 > gpurun_out/g4/tests_multi.log 2>&1; tail -25 gpurun_out/g4/tests_multi.log
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline --no-prefill "$@" > gpurun_out/g4/bench_$tag.json 2> gpurun_out/g4/bench_$tag.err; tail -c 900 gpurun_out/g4/bench_$tag.json; tail -3 gpurun_out/g4/bench_$tag.err; }
run tp2
run tp2_slabs1_nograph --slabs 1 --no-graph
run tp2_slabs4 --slabs 4
run tp2_ar --parallelism tp-ar
run tp2_70b --workload llama3-70b
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-prefill > gpurun_out/g4/bench_n1.json 2> gpurun_out/g4/bench_n1.err; tail -c 600 gpurun_out/g4/bench_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/prefill_bench.py --model qwen2.5-7b --bsz 4 --seq 1024 > gpurun_out/g4/prefill_qwen7b_tp2.json 2> gpurun_out/g4/prefill_tp2.err; cat gpurun_out/g4/prefill_qwen7b_tp2.json; tail -3 gpurun_out/g4/prefill_tp2.err
