# tile-organisation sweep: bash scripts/quant_sweep.sh "cfg1" "cfg2" ...   (cfg = ARCQ_QUANT_TMA knob, "-" = default dispatch)
for cfg in "$@"; do
  echo "cfg $cfg"
  if [ "$cfg" = "-" ]; then unset ARCQ_QUANT_TMA; else export ARCQ_QUANT_TMA=$cfg; fi
  timeout 100 python scripts/quant_prof.py --shapes 32768,4096,256 8192,4096,256 8192,3584,256 2>&1 | grep quant
done
