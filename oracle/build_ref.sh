#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds the reference's own sm80 CuTe kernel TU for sm_100a,
# straight from the sources where they lie under /root/reference (nothing is copied into
# this repo), plus our C-ABI shim oracle/ref_shim.cu -> oracle/_ref/libref_fa.so.
# Flags follow the reference loader (flash_attention/load_cpp_extention.py:30-51).
# The output is git-ignored but travels to the GPU box with the gpurun snapshot.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF_ROOT:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -f "$REF/csrc/flash_attention_impl.cu" ]; then
  echo "reference sources not present at $REF; keeping any prebuilt $OUT/libref_fa.so" >&2
  exit 0
fi
mkdir -p "$OUT"
if [ -f "$OUT/libref_fa.so" ] && [ "$OUT/libref_fa.so" -nt "$HERE/ref_shim.cu" ] && [ -z "${FORCE:-}" ]; then
  echo "up to date: $OUT/libref_fa.so"; exit 0
fi
nvcc -O3 -std=c++17 --expt-relaxed-constexpr --expt-extended-lambda --use_fast_math -lineinfo \
  -gencode arch=compute_100a,code=sm_100a \
  -I "$REF/csrc" -I "$REF/csrc/3rd/cutlass/include" -I "$REF/csrc/3rd/cutlass/tools/util/include" \
  -shared -Xcompiler -fPIC -o "$OUT/libref_fa.so" \
  "$HERE/ref_shim.cu" "$REF/csrc/flash_attention_impl.cu"
echo "built $OUT/libref_fa.so"
