#!/usr/bin/env bash
# build_ref.sh -- TEST INFRASTRUCTURE ONLY.  Compiles, FROM WHERE THEY LIE under /root/reference, (a) a host
# probe of the reference's arithmetic library (CUTLASS converters + CuTe SF layout) and (b) the reference's own
# quantise kernels, unchanged, for sm_100a, behind oracle/refkernels_shim.cu.  Outputs only into oracle/_ref/
# (git-ignored; travels to the GPU box with the gpurun snapshot).  No reference source is copied.
# (c) "py": the reference's own PYTHON CALLERS of the extension (model/qLinearLayer.py + the model/quantize.py it imports,
# kernels/main.py, benchmarks/modeling_arc.py) copied UNMODIFIED into the git-ignored oracle/_ref/py/ so that the -m gpu
# test tests/test_gpu_ref_callers.py can run them on top of THIS repo's `agemm` on the GPU box, where /root/reference does
# not exist.  They never enter the repository's history (oracle/_ref/ is in .gitignore).
# (d) "gemm": oracle/ref_cutlass_gemm.cu = the reference's CUTLASS block-scaled GEMM instantiation (kernels/src/nvfp4.cu:28-74)
# retargeted Sm120 -> Sm100 with the tile / cluster of CUTLASS example 72a, against the reference's vendored CUTLASS headers
# -> oracle/_ref/libarcq_refgemm.so: a device-side GEMM oracle and the "reference config retargeted" number to beat (~3 min).
#   usage: oracle/build_ref.sh [probe|kernels|py|gemm|all]   (default: all; kernels take ~4 min of nvcc)
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference
OUT="$HERE/_ref"
WHAT="${1:-all}"
[ -d "$REF/kernels/src" ] || { echo "build_ref: $REF not present (GPU box?) -- using prebuilt $OUT"; exit 0; }
mkdir -p "$OUT/tables"
INC="-I$REF/kernels/include -I$REF/third-party/cutlass/include -I$REF/third-party/cutlass/tools/util/include -I$REF/third-party/cutlass/examples/common"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
if [ "$WHAT" = probe ] || [ "$WHAT" = all ]; then
  if [ ! -x "$OUT/ref_probe" ] || [ "$HERE/ref_probe.cu" -nt "$OUT/ref_probe" ]; then
    $NVCC -std=c++17 -O1 -w --expt-relaxed-constexpr -arch=sm_100a $INC "$HERE/ref_probe.cu" -o "$OUT/ref_probe"
  fi
  "$OUT/ref_probe" "$OUT/tables"
fi
if [ "$WHAT" = py ] || [ "$WHAT" = all ]; then
  mkdir -p "$OUT/py"
  cp "$REF/model/qLinearLayer.py" "$REF/model/quantize.py" "$REF/kernels/main.py" "$REF/benchmarks/modeling_arc.py" "$OUT/py/"
  echo "build_ref: copied the reference callers into $OUT/py (git-ignored)"
fi
if [ "$WHAT" = gemm ] || [ "$WHAT" = all ]; then
  if [ ! -f "$OUT/libarcq_refgemm.so" ] || [ "$HERE/ref_cutlass_gemm.cu" -nt "$OUT/libarcq_refgemm.so" ]; then
    $NVCC -std=c++17 -O2 -w --expt-relaxed-constexpr -Xcompiler -fPIC -gencode arch=compute_100a,code=sm_100a \
      -DCUTLASS_ARCH_MMA_SM100_SUPPORTED=1 -I"$REF/third-party/cutlass/include" -I"$REF/third-party/cutlass/tools/util/include" \
      -shared -o "$OUT/libarcq_refgemm.so" "$HERE/ref_cutlass_gemm.cu" &
    GEMM_PID=$!
  fi
fi
if [ "$WHAT" = kernels ] || [ "$WHAT" = all ]; then
  FLAGS="-std=c++17 -O2 -w --expt-relaxed-constexpr -Xcompiler -fPIC -Xcompiler -fpermissive -gencode arch=compute_100a,code=sm_100a $INC"
  pids=()
  for f in reorder down rmsnorm; do
    if [ ! -f "$OUT/$f.o" ] || [ "$REF/kernels/src/$f.cu" -nt "$OUT/$f.o" ]; then
      $NVCC $FLAGS -c "$REF/kernels/src/$f.cu" -o "$OUT/$f.o" & pids+=($!)
    fi
  done
  $NVCC $FLAGS -c "$HERE/refkernels_shim.cu" -o "$OUT/shim.o" & pids+=($!)
  for p in "${pids[@]}"; do wait "$p"; done
  $NVCC -shared -Wno-deprecated-gpu-targets -o "$OUT/libarcq_refkernels.so" "$OUT/reorder.o" "$OUT/down.o" "$OUT/rmsnorm.o" "$OUT/shim.o"
  echo "build_ref: built $OUT/libarcq_refkernels.so"
fi
if [ -n "${GEMM_PID:-}" ]; then wait "$GEMM_PID" && echo "build_ref: built $OUT/libarcq_refgemm.so"; fi
