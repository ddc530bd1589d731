for cfg in "4,256,2,4,2" "4,256,1,4,3" "4,256,1,4,2" "2,256,2,2,3" "2,256,2,2,4" "2,256,3,2,3" "2,256,4,2,2" "2,256,1,2,4" "4,512,2,4,1" "2,512,2,2,1"; do
  echo "cfg $cfg"; ARCQ_QUANT_TMA=$cfg timeout 100 python scripts/quant_prof.py --shapes 32768,4096,256 8192,4096,256 2>&1 | grep quant
done
