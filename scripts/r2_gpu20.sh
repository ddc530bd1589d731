#!/bin/bash
# GPU call 20 (1 GPU): grouped tile order of the pair GEMM -- parity, then A/B inside the bench step
mkdir -p gpurun_out/g20
timeout 200 python -m pytest tests/test_gpu_gemm_large.py -q -x -p no:cacheprovider -k "grouped or default_dispatch" > gpurun_out/g20/tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/g20/tests.log
for g in 0 8 4 16; do
  timeout 100 python bench.py --steps 20 --warmup 5 --no-prefill --no-cpu-baseline --e2e-steps 1 --no-check --option gemm_raster=$g > gpurun_out/g20/bench_g$g.json 2> gpurun_out/g20/bench_g$g.err
  python - $g <<'PY'
import json,sys
g=sys.argv[1]
o=json.loads([l for l in open("gpurun_out/g20/bench_g%s.json"%g) if l.startswith("{")][-1])
print("raster",g, round(o["ms_per_step"],4), [(k["launch"], round(k["us"],1)) for k in o["kernels"] if k["kernel"]=="gemm"])
PY
done
