#!/bin/bash
# GPU call 13 (1 GPU): fused-add tests after the revert, split-K over several row tiles (tests + A/B)
mkdir -p gpurun_out/g13
timeout 600 python -m pytest tests/test_gpu_addnorm.py tests/test_gpu_gemm_large.py tests/test_gpu_gemm.py tests/test_gpu_quant_tiles.py -q -x -p no:cacheprovider > gpurun_out/g13/tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/g13/tests.log
timeout 600 python scripts/midm_ab.py > gpurun_out/g13/midm_ab.log 2>&1; cat gpurun_out/g13/midm_ab.log
