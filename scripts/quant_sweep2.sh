# usage: bash scripts/quant_sweep2.sh "<quant_prof args>" cfg1 cfg2 ...
args="$1"; shift
for cfg in "$@"; do
  echo "cfg $cfg"
  if [ "$cfg" = "-" ]; then unset ARCQ_QUANT_TMA; else export ARCQ_QUANT_TMA=$cfg; fi
  timeout 100 python scripts/quant_prof.py $args 2>&1 | grep quant
done
