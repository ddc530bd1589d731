#!/bin/bash
# round-2 final validation on ONE B200: full GPU suite, smoke, bench (both arms), launch list, per-kernel ncu capture
mkdir -p gpurun_out/final2
timeout 400 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/final2/tests_gpu.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/final2/tests_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/final2/smoke.log 2>&1; tail -1 gpurun_out/final2/smoke.log
timeout 200 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/final2/bench_ref.json 2> gpurun_out/final2/bench_ref.err; tail -c 400 gpurun_out/final2/bench_ref.json
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/final2/bench.json 2> gpurun_out/final2/bench.err; echo "bench rc=$?"; python - <<'PY'
import json
o=json.loads([l for l in open("gpurun_out/final2/bench.json") if l.startswith("{")][-1])
print({k:o[k] for k in ("value","ms_per_step","gpu_launches","binding")}, o["e2e"], o["roofline"]["frac"], o["roofline_quant"]["frac"], o["parity_check"]["ok"], o["prefill_config3"], o["clocks"])
PY
tail -3 gpurun_out/final2/bench.err
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final2/launches.csv python bench.py --steps 2 --warmup 1 --e2e-steps 1 --no-prefill --no-cpu-baseline > gpurun_out/final2/ncu_bench.log 2>&1; wc -l gpurun_out/final2/launches.csv
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'quant|gemm' -s 12 -c 6 -o gpurun_out/final2/step python scripts/step_prof.py > gpurun_out/final2/ncu_step.log 2>&1; tail -1 gpurun_out/final2/ncu_step.log
