# round 2, GPU call 3
set -x
mkdir -p gpurun_out/g3
timeout 2400 python -m pytest tests -q -m gpu --maxfail=40 -p no:cacheprovider --timeout 600 > gpurun_out/g3/tests_gpu.log 2>&1; tail -12 gpurun_out/g3/tests_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/g3/bench.json 2> gpurun_out/g3/bench.err; tail -c 2500 gpurun_out/g3/bench.json; tail -5 gpurun_out/g3/bench.err
timeout 600 python bench.py --steps 20 --warmup 5 --unfused --no-prefill --no-cpu-baseline > gpurun_out/g3/bench_unfused.json 2> gpurun_out/g3/bench_unfused.err; tail -c 600 gpurun_out/g3/bench_unfused.json
timeout 900 python scripts/measure_nvfp4_peak.py > gpurun_out/g3/r2_nvfp4_peak.json 2> gpurun_out/g3/peak.err; tail -5 gpurun_out/g3/peak.err
timeout 900 python scripts/config2_sweep.py > gpurun_out/g3/r2_sweep.json 2> gpurun_out/g3/sweep.err; tail -5 gpurun_out/g3/sweep.err; tail -c 600 gpurun_out/g3/r2_sweep.json
timeout 300 python scripts/quant_prof.py > gpurun_out/g3/quant_plan.log 2>&1; cat gpurun_out/g3/quant_plan.log
