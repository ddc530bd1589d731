# round 2, GPU call 6 (1 GPU): full test pass, default bench (both arms), ncu launch list + --set full of the step
set -x
mkdir -p gpurun_out/g6
timeout 600 python -m pytest tests -q -m gpu --maxfail=20 -p no:cacheprovider --timeout 300 > gpurun_out/g6/tests_gpu.log 2>&1; tail -4 gpurun_out/g6/tests_gpu.log | cut -c1-300
timeout 120 python __graft_entry__.py smoke > gpurun_out/g6/smoke.log 2>&1; tail -2 gpurun_out/g6/smoke.log
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/g6/bench_ref.json 2> gpurun_out/g6/bench_ref.err; tail -c 300 gpurun_out/g6/bench_ref.json
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/g6/bench.json 2> gpurun_out/g6/bench.err; tail -c 600 gpurun_out/g6/bench.json; tail -3 gpurun_out/g6/bench.err
timeout 120 python scripts/step_prof.py > gpurun_out/g6/step_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'quant|gemm' -s 12 -c 6 -o gpurun_out/g6/step python scripts/step_prof.py > gpurun_out/g6/ncu_step.log 2>&1; tail -3 gpurun_out/g6/ncu_step.log
timeout 200 python bench.py --steps 2 --warmup 1 --e2e-steps 1 --no-cpu-baseline --no-prefill --no-check > gpurun_out/g6/bench_short.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/g6/launches.csv python bench.py --steps 2 --warmup 1 --e2e-steps 1 --no-cpu-baseline --no-prefill --no-check > gpurun_out/g6/ncu_bench.log 2>&1; tail -2 gpurun_out/g6/ncu_bench.log | cut -c1-200
ls -la gpurun_out/g6
