# round 2, GPU call 8 (1 GPU): tile-organisation A/B with plans
set -x
mkdir -p gpurun_out/g8
S="--shapes 8192,4096,256 32768,4096,256 8192,3584,256 8192,2048,128"
timeout 120 python scripts/quant_prof.py $S > gpurun_out/g8/q_default.log 2>&1; cat gpurun_out/g8/q_default.log
timeout 120 python scripts/quant_prof.py $S --tile 4,256,2,4,2 > gpurun_out/g8/q_r4x2_s2.log 2>&1; cat gpurun_out/g8/q_r4x2_s2.log
timeout 120 python scripts/quant_prof.py $S --tile 4,256,1,4,2 > gpurun_out/g8/q_r4x2_s1.log 2>&1; cat gpurun_out/g8/q_r4x2_s1.log
timeout 120 python scripts/quant_prof.py $S --tile 8,512,1,8,1 > gpurun_out/g8/q_r8rt8.log 2>&1; cat gpurun_out/g8/q_r8rt8.log
timeout 120 python scripts/quant_prof.py --shapes 8192,8192,512 32768,8192,512 --tile 4,512,2,4,1 > gpurun_out/g8/q_8192_s2.log 2>&1; cat gpurun_out/g8/q_8192_s2.log
timeout 120 python scripts/quant_prof.py --shapes 8192,8192,512 32768,8192,512 > gpurun_out/g8/q_8192_def.log 2>&1; cat gpurun_out/g8/q_8192_def.log
