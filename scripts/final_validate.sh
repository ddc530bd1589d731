# round-end validation on one B200: tests, smoke, bench (both arms), launch list, per-kernel ncu capture, quantise table
set -x
mkdir -p gpurun_out/final
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/final/tests_gpu.log 2>&1; tail -3 gpurun_out/final/tests_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/final/smoke.log 2>&1; tail -2 gpurun_out/final/smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/final/bench_ref.json 2> gpurun_out/final/bench_ref.err; tail -c 600 gpurun_out/final/bench_ref.json
timeout 600 python bench.py > gpurun_out/final/bench.json 2> gpurun_out/final/bench.err; tail -c 1500 gpurun_out/final/bench.json
timeout 300 python scripts/quant_prof.py > gpurun_out/final/quant.log 2>&1
timeout 300 python scripts/quant_prof.py --rms --shapes 8192,4096,256 32768,4096,256 8192,3584,256 8192,8192,512 >> gpurun_out/final/quant.log 2>&1
cat gpurun_out/final/quant.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final/launches.csv python bench.py --steps 2 --warmup 1 --e2e-steps 1 > gpurun_out/final/ncu_bench.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'quant|gemm' -s 16 -c 6 -o gpurun_out/final/step python scripts/step_prof.py > gpurun_out/final/ncu_step.log 2>&1; tail -2 gpurun_out/final/ncu_step.log
