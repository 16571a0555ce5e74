#!/usr/bin/env bash
# Compiles the subset of the reference that builds unmodified against pyarrow's Arrow C++ (24.x) into
# oracle/_ref/libcylon_ref.so: hash partition kernels, row comparators, the hash join, murmur3. Sources are compiled
# IN PLACE from /root/reference (never copied). The reference's own build system (cmake + Arrow 2.0 + MPI + glog
# fetched from the network) cannot run here; see DESIGN.md for what does not compile and why.
set -euo pipefail
OUT="$1"
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference/cpp/src/cylon
PA_INC=$(python -c "import pyarrow as pa; print(pa.get_include())")
PA_LIB=$(python -c "import pyarrow as pa; print(pa.get_library_dirs()[0])")
LIBARROW=$(ls "$PA_LIB"/libarrow.so.* | head -1)
LIBARROW_COMPUTE=$(ls "$PA_LIB"/libarrow_compute.so.* | head -1)
SRCS="$REF/join/join.cpp $REF/join/hash_join.cpp $REF/join/sort_join.cpp $REF/join/join_utils.cpp \
      $REF/arrow/arrow_partition_kernels.cpp $REF/arrow/arrow_comparator.cpp $REF/util/arrow_utils.cpp $REF/arrow/arrow_kernels.cpp \
      $REF/util/murmur3.cpp"
if [ -f "$OUT/libcylon_ref.so" ] && [ "$OUT/libcylon_ref.so" -nt "$HERE/cylon_ref_shim.cpp" ]; then exit 0; fi
g++ -std=c++20 -O3 -fPIC -shared -w -include "$HERE/stubs/compat.h" -I"$HERE/stubs" -I"$REF" -I"$PA_INC" -o "$OUT/libcylon_ref.so" \
    "$HERE/cylon_ref_shim.cpp" $SRCS "$LIBARROW_COMPUTE" "$LIBARROW" -Wl,-rpath,"$PA_LIB"
echo "build_cylon_ref: built $OUT/libcylon_ref.so"
