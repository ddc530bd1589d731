# round 2, first GPU call: parity of everything new, peaks, first bench lines (one GPU)
set -x
mkdir -p gpurun_out/g1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -c "from arcquant_b200 import agemm; print('binding', agemm.BACKEND)"
timeout 2400 python -m pytest tests -q -m gpu --maxfail=40 -p no:cacheprovider --timeout 600 > gpurun_out/g1/tests_gpu.log 2>&1; tail -60 gpurun_out/g1/tests_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/g1/smoke.log 2>&1; tail -3 gpurun_out/g1/smoke.log
timeout 900 python scripts/measure_nvfp4_peak.py > gpurun_out/g1/r2_nvfp4_peak.json 2> gpurun_out/g1/peak.err; tail -c 1500 gpurun_out/g1/r2_nvfp4_peak.json; tail -5 gpurun_out/g1/peak.err
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/g1/bench.json 2> gpurun_out/g1/bench.err; tail -c 3000 gpurun_out/g1/bench.json; tail -5 gpurun_out/g1/bench.err
timeout 600 python bench.py --steps 20 --warmup 5 --unfused --no-prefill --no-cpu-baseline > gpurun_out/g1/bench_unfused.json 2> gpurun_out/g1/bench_unfused.err; tail -c 1500 gpurun_out/g1/bench_unfused.json
timeout 300 python scripts/quant_prof.py > gpurun_out/g1/quant.log 2>&1; cat gpurun_out/g1/quant.log
