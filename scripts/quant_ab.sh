# same-box A/B of library variants: bash scripts/quant_ab.sh name1 name2 ...   (b200 = the default build)
for v in "$@"; do
  for rep in 1 2; do
    echo "== variant $v rep $rep"
    ARCQ_LIB=arcquant_b200/libarcq_$v.so timeout 200 python scripts/quant_prof.py --shapes 8192,4096,256 32768,4096,256 8192,14336,512 8192,8192,512 2>&1 | grep quant
    ARCQ_LIB=arcquant_b200/libarcq_$v.so timeout 200 python scripts/quant_prof.py --rms --shapes 8192,4096,256 32768,4096,256 8192,8192,512 2>&1 | grep quant
  done
done
