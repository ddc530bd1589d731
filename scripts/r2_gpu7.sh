# round 2, GPU call 7 (8 GPUs, strictly bounded): scaling lines for config 2/4 and the config-5 prefill
set -x
mkdir -p gpurun_out/g7
run() { n=$1; tag=$2; shift 2; timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline --no-prefill "$@" > gpurun_out/g7/bench_$tag.json 2> gpurun_out/g7/bench_$tag.err; echo "rc=$?"; tail -c 300 gpurun_out/g7/bench_$tag.json; tail -2 gpurun_out/g7/bench_$tag.err | cut -c1-200; }
run 8 tp8
run 4 tp4
run 8 tp8_70b --workload llama3-70b
run 4 tp4_70b --workload llama3-70b
run 8 tp8_ar --parallelism tp-ar
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 scripts/prefill_bench.py --model qwen2.5-32b --bsz 8 --seq 2048 > gpurun_out/g7/prefill_qwen32b_tp8.json 2> gpurun_out/g7/prefill_tp8.err; cat gpurun_out/g7/prefill_qwen32b_tp8.json | cut -c1-500; tail -2 gpurun_out/g7/prefill_tp8.err | cut -c1-300
