#!/bin/bash
# ORACLE -- TEST INFRASTRUCTURE ONLY.
# Compiles the UNMODIFIED reference sources where they lie (/root/reference/src) for sm_100,
# plus oracle/ref_harness.cu, into oracle/_ref/libref_tcqr.so (git-ignored, travels to the GPU
# box with the snapshot). Also dumps the reference's exported mtk:: symbols, which
# tests/golden/reference_symbols.txt pins (the drop-in link contract).
# Flags follow the reference Makefile:3-6 except the gencode (its list stops at sm_86).
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
[ -d "$REF/src" ] || { echo "reference tree not present at $REF (GPU box): keeping prebuilt $OUT"; exit 0; }
mkdir -p "$OUT"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-std=c++14 -O3 -rdc=true --expt-relaxed-constexpr -Xcompiler=-fPIC -Xcompiler=-fopenmp -w \
  -gencode arch=compute_100,code=sm_100 \
  -I $REF/src/cutf -I $REF/src/wmma_extension/include -I $REF/src/runtime_status -I $REF/src"
pids=()
for f in tcqr32x16 tsqr blockqr; do
  $NVCC $FLAGS -c "$REF/src/$f.cu" -o "$OUT/$f.o" & pids+=($!)
done
$NVCC $FLAGS -c "$HERE/ref_harness.cu" -o "$OUT/ref_harness.o" & pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done
$NVCC -shared -gencode arch=compute_100,code=sm_100 -Xlinker -Bsymbolic -o "$OUT/libref_tcqr.so" \
  "$OUT"/tcqr32x16.o "$OUT"/tsqr.o "$OUT"/blockqr.o "$OUT"/ref_harness.o -lcublas -lcudart -lgomp
nm --defined-only "$OUT"/tcqr32x16.o "$OUT"/tsqr.o "$OUT"/blockqr.o | awk '$2 ~ /^[TW]$/ && $3 ~ /^_ZN3mtk/ {print $3}' \
  | grep -E '^_ZN3mtk(2qr|4tsqr|4tcqr)' | sort -u > "$OUT/reference_symbols.txt"
echo "built $OUT/libref_tcqr.so; $(wc -l < "$OUT/reference_symbols.txt") reference symbols"
