# round 2, GPU call 9 (1 GPU): no-alloc entry point alone in a fresh process, full GPU suite, smoke, bench, quantise tile A/B
set -x
mkdir -p gpurun_out/g9
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -p no:cacheprovider --timeout 200 > gpurun_out/g9/noalloc_alone.log 2>&1; tail -5 gpurun_out/g9/noalloc_alone.log
timeout 1100 python -m pytest tests -q -m gpu --maxfail=20 -p no:cacheprovider --timeout 300 > gpurun_out/g9/pytest_gpu.log 2>&1; tail -25 gpurun_out/g9/pytest_gpu.log
timeout 200 python __graft_entry__.py smoke > gpurun_out/g9/smoke.log 2>&1; tail -3 gpurun_out/g9/smoke.log
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/g9/bench.json 2> gpurun_out/g9/bench.err; cat gpurun_out/g9/bench.json; tail -3 gpurun_out/g9/bench.err
S="--shapes 8192,4096,256 32768,4096,256 8192,3584,256 8192,2048,128 8192,8192,512 8192,14336,512 32768,14336,512"
timeout 150 python scripts/quant_prof.py $S > gpurun_out/g9/q_default.log 2>&1; cat gpurun_out/g9/q_default.log
timeout 150 python scripts/quant_prof.py $S --tile 8,512,1,4,1 > gpurun_out/g9/q_r8rt4.log 2>&1; cat gpurun_out/g9/q_r8rt4.log
timeout 150 python scripts/quant_prof.py $S --no-plans > gpurun_out/g9/q_noplans.log 2>&1; cat gpurun_out/g9/q_noplans.log
