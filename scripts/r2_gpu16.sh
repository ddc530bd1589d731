#!/bin/bash
# GPU call 16 (1 GPU): bias epilogue tests, prefill with it, refreshed config-2 sweep, ncu of the fused add kernel
mkdir -p gpurun_out/g16
timeout 600 python -m pytest tests/test_gpu_gemm_large.py tests/test_gpu_gemm.py tests/test_gpu_prefill.py tests/test_gpu_tp.py tests/test_gpu_ref_callers.py -q -x -p no:cacheprovider > gpurun_out/g16/tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/g16/tests.log
timeout 300 python scripts/prefill_bench.py > gpurun_out/g16/prefill.log 2>&1; tail -1 gpurun_out/g16/prefill.log
timeout 600 python scripts/config2_sweep.py > gpurun_out/g16/sweep.json 2> gpurun_out/g16/sweep.err; tail -2 gpurun_out/g16/sweep.err; wc -c gpurun_out/g16/sweep.json
timeout 300 ncu --set full --clock-control none --import-source on -k regex:quant_tma -s 4 -c 1 -o gpurun_out/g16/addnorm python scripts/quant_prof.py --rms --add --shapes 8192,4096,256 --iters 2 > gpurun_out/g16/ncu_add.log 2>&1; tail -2 gpurun_out/g16/ncu_add.log
