# round 2, GPU call 2: fixed tests, plan A/B, fused-launch bench, ncu of the step
set -x
mkdir -p gpurun_out/g2
timeout 2400 python -m pytest tests -q -m gpu --maxfail=40 -p no:cacheprovider --timeout 600 > gpurun_out/g2/tests_gpu.log 2>&1; tail -40 gpurun_out/g2/tests_gpu.log
timeout 300 python scripts/quant_prof.py > gpurun_out/g2/quant_plan.log 2>&1; cat gpurun_out/g2/quant_plan.log
timeout 300 python scripts/quant_prof.py --no-plans > gpurun_out/g2/quant_noplan.log 2>&1; cat gpurun_out/g2/quant_noplan.log
timeout 300 python scripts/quant_prof.py --shapes 32768,4096,256 32768,14336,512 32768,8192,512 > gpurun_out/g2/quant_plan_32k.log 2>&1; cat gpurun_out/g2/quant_plan_32k.log
timeout 300 python scripts/quant_prof.py --no-plans --shapes 32768,4096,256 32768,14336,512 32768,8192,512 > gpurun_out/g2/quant_noplan_32k.log 2>&1; cat gpurun_out/g2/quant_noplan_32k.log
timeout 300 python scripts/quant_prof.py --rms --shapes 8192,4096,256 32768,4096,256 8192,3584,256 8192,8192,512 > gpurun_out/g2/quant_rms.log 2>&1; cat gpurun_out/g2/quant_rms.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/g2/bench.json 2> gpurun_out/g2/bench.err; tail -c 2500 gpurun_out/g2/bench.json; tail -5 gpurun_out/g2/bench.err
timeout 600 python bench.py --steps 20 --warmup 5 --unfused --no-prefill --no-cpu-baseline > gpurun_out/g2/bench_unfused.json 2> gpurun_out/g2/bench_unfused.err; tail -c 600 gpurun_out/g2/bench_unfused.json
timeout 900 python scripts/measure_nvfp4_peak.py > gpurun_out/g2/r2_nvfp4_peak.json 2> gpurun_out/g2/peak.err; tail -5 gpurun_out/g2/peak.err
