#!/bin/bash
# GPU call 12 (1 GPU): rmsnorm on the planned path, fused residual add, prefill with it
mkdir -p gpurun_out/g12
timeout 500 python -m pytest tests/test_gpu_addnorm.py tests/test_gpu_plan.py tests/test_gpu_quant_tiles.py tests/test_gpu_refkernels.py -q -x -p no:cacheprovider > gpurun_out/g12/tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/g12/tests.log
timeout 300 python -m pytest tests/test_gpu_prefill.py tests/test_gpu_peer.py -q -x -p no:cacheprovider > gpurun_out/g12/tests2.log 2>&1; echo "tests2 rc=$?"; tail -5 gpurun_out/g12/tests2.log
S="--shapes 8192,4096,256 32768,4096,256 8192,3584,256 8192,8192,512 8192,2048,128"
timeout 150 python scripts/quant_prof.py $S --rms > gpurun_out/g12/q_rms_plan.log 2>&1; cat gpurun_out/g12/q_rms_plan.log
timeout 150 python scripts/quant_prof.py $S --rms --no-plans > gpurun_out/g12/q_rms_noplan.log 2>&1; cat gpurun_out/g12/q_rms_noplan.log
timeout 150 python scripts/quant_prof.py $S --rms --tile 8,512,1,4,1 > gpurun_out/g12/q_rms_plan_r8.log 2>&1; cat gpurun_out/g12/q_rms_plan_r8.log
timeout 150 python scripts/quant_prof.py $S --rms --add > gpurun_out/g12/q_rms_add.log 2>&1; cat gpurun_out/g12/q_rms_add.log
timeout 150 python scripts/quant_prof.py $S > gpurun_out/g12/q_plain.log 2>&1; cat gpurun_out/g12/q_plain.log
timeout 300 python scripts/prefill_bench.py > gpurun_out/g12/prefill.log 2>&1; tail -3 gpurun_out/g12/prefill.log
