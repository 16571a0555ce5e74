#!/usr/bin/env bash
# Builds checker libraries from the reference's OWN sources, in place under /root/reference
# (nothing is copied into this repo). Outputs go to oracle/_ref/ (git-ignored, travels with gpurun).
# Skipped silently when /root/reference is absent (the GPU box uses the prebuilt files).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference/cpp/src/cylon
OUT="$HERE/_ref"
[ -d "$REF" ] || { echo "build_ref: $REF absent, keeping prebuilt files"; exit 0; }
mkdir -p "$OUT"
CXX=${CXX:-g++}
# 1. murmur3 (Arrow-free)
if [ ! -f "$OUT/libref_murmur3.so" ] || [ "$HERE/ref_shim/murmur_shim.cpp" -nt "$OUT/libref_murmur3.so" ]; then
  $CXX -O2 -fPIC -shared -std=c++17 -I"$REF" -o "$OUT/libref_murmur3.so" \
      "$HERE/ref_shim/murmur_shim.cpp" "$REF/util/murmur3.cpp"
fi
# 2. join / partition / set-op kernels against pyarrow's Arrow (see ref_shim/README.md); optional
if [ -x "$HERE/ref_shim/build_cylon_ref.sh" ]; then
  "$HERE/ref_shim/build_cylon_ref.sh" "$OUT" || echo "build_ref: reference join library not built (see DESIGN.md)"
fi
# 3. aggregation kernels (header-only in the reference: compute/aggregate_kernels.hpp), Arrow-free
if [ ! -f "$OUT/libref_aggregate.so" ] || [ "$HERE/ref_shim/aggregate_shim.cpp" -nt "$OUT/libref_aggregate.so" ]; then
  $CXX -O2 -fPIC -shared -std=c++17 -I"$REF" -o "$OUT/libref_aggregate.so" "$HERE/ref_shim/aggregate_shim.cpp"
fi
echo "build_ref: ok"
